/* abi_harness.c -- a plain C11 caller of include/bsx_b200.h: what a non-Python host (the Rust shim's FFI, a C++
 * prover) does through the drop-in boundary.  No Python, no torch; links libbsx_b200.so only.
 *
 *   abi_harness <log_n> <n_cols> <rate_bits> <cap_height> <seed>
 *
 * Commits a SplitMix64 witness from host memory, then prints, one record per line (hex words):
 *   cap <2^cap_height x 4 words>
 *   row <leaf> <n_cols words>          for three leaves
 *   path <leaf> <siblings x 4 words>
 *   eval <n_cols x 2 words>            all polynomials at the extension point (seed, seed + 1)
 *   proof <words>                      prove_openings of every polynomial at that point
 *   comm <G> <same_cap> <same_rows_paths> <same_proof>   the same commit and proof through bsx_comm_create with G ranks
 *                                      emulated on device 0 (bsx_sharded_from_values / bsx_sharded_prove_openings)
 *   fribytes <len> <fnv64>             bsx_fri_proof_bytes of that proof; layout <total_words> from bsx_fri_proof_layout
 *   qv <log_n> <fnv64> <first> <last>  bsx_quotient_values of a fourteen-kind gate table over three commitments of SplitMix columns
 *   pgw <2 x 135 words>                bsx_poseidon_gate_witness on two rows (swap 0 and 1)
 * tests/test_gpu_c_abi.py recomputes every record with the oracle. */
#include <inttypes.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "bsx_b200.h"

#define P 0xFFFFFFFF00000001ULL

static uint64_t splitmix(uint64_t seed, uint64_t i) {
    uint64_t z = seed + (i + 1) * 0x9E3779B97F4A7C15ULL;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    z ^= z >> 31;
    return z % P;
}

#define CHECK(ctx, call)                                                                 \
    do {                                                                                 \
        int32_t rc_ = (call);                                                            \
        if (rc_ != BSX_OK) {                                                             \
            fprintf(stderr, "%s -> %d: %s\n", #call, (int)rc_, bsx_last_error(ctx));     \
            return 1;                                                                    \
        }                                                                                \
    } while (0)

static void print_words(const char *tag, long long k, const uint64_t *w, size_t n) {
    if (k >= 0) printf("%s %lld", tag, k);
    else printf("%s", tag);
    for (size_t i = 0; i < n; i++) printf(" %016" PRIx64, w[i]);
    printf("\n");
}

int main(int argc, char **argv) {
    if (argc != 6) {
        fprintf(stderr, "usage: %s log_n n_cols rate_bits cap_height seed\n", argv[0]);
        return 2;
    }
    const int log_n = atoi(argv[1]), n_cols = atoi(argv[2]), rate_bits = atoi(argv[3]), cap_height = atoi(argv[4]);
    const uint64_t seed = strtoull(argv[5], NULL, 0);
    const size_t n = (size_t)1 << log_n, nb = n << rate_bits;

    bsx_ctx *ctx = NULL;
    if (bsx_ctx_create(0, &ctx) != BSX_OK) {
        fprintf(stderr, "bsx_ctx_create failed (no CUDA device? there is no CPU fallback)\n");
        return 1;
    }
    uint64_t *values = malloc(sizeof(uint64_t) * n * (size_t)n_cols);
    for (size_t i = 0; i < n * (size_t)n_cols; i++) values[i] = splitmix(seed, i);

    bsx_batch *batch = NULL;
    bsx_timing tm;
    CHECK(ctx, bsx_batch_from_values(ctx, values, BSX_HOST, log_n, n_cols, rate_bits, cap_height, 0, NULL, &batch, &tm));
    free(values);

    uint64_t *cap = malloc(sizeof(uint64_t) * ((size_t)4 << cap_height));
    CHECK(ctx, bsx_batch_cap(batch, cap));
    print_words("cap", -1, cap, (size_t)4 << cap_height);

    bsx_tree *tree = bsx_batch_tree(batch);
    const int n_sib = log_n + rate_bits - cap_height;
    uint64_t *row = malloc(sizeof(uint64_t) * (size_t)n_cols), *sib = malloc(sizeof(uint64_t) * 4 * (size_t)(n_sib + 1));
    const uint64_t leaves[3] = {0, nb / 3, nb - 1};
    for (int k = 0; k < 3; k++) {
        CHECK(ctx, bsx_merkle_get(tree, &leaves[k], 1, row));
        print_words("row", (long long)leaves[k], row, (size_t)n_cols);
        CHECK(ctx, bsx_merkle_prove(tree, leaves[k], sib));
        print_words("path", (long long)leaves[k], sib, 4 * (size_t)n_sib);
    }

    const uint64_t z[2] = {seed % P, (seed + 1) % P};
    uint64_t *ev = malloc(sizeof(uint64_t) * 2 * (size_t)n_cols);
    CHECK(ctx, bsx_batch_eval_ext(batch, 0, n_cols, z, 1, ev));
    print_words("eval", -1, ev, 2 * (size_t)n_cols);

    /* prove_openings: one opening batch holding every polynomial, standard FRI shape scaled to this size */
    bsx_fri_params fp;
    memset(&fp, 0, sizeof(fp));
    fp.degree_bits = log_n;
    fp.rate_bits = rate_bits;
    fp.cap_height = cap_height;
    fp.proof_of_work_bits = 8;
    fp.num_query_rounds = 4;
    int db = log_n;
    while (db > 5 && db + rate_bits - 4 >= cap_height && fp.n_arity < 16) { /* ConstantArityBits(4, 5) */
        fp.arity_bits[fp.n_arity++] = 4;
        db -= 4;
    }
    bsx_fri_instance inst;
    memset(&inst, 0, sizeof(inst));
    int32_t *po = malloc(sizeof(int32_t) * (size_t)n_cols), *pi = malloc(sizeof(int32_t) * (size_t)n_cols);
    for (int j = 0; j < n_cols; j++) {
        po[j] = 0;
        pi[j] = j;
    }
    int32_t n_polys[1] = {n_cols};
    inst.n_batches = 1;
    inst.points = z;
    inst.n_polys = n_polys;
    inst.poly_oracle = po;
    inst.poly_index = pi;
    bsx_challenger ch;
    bsx_challenger_init(&ch);
    CHECK(ctx, bsx_challenger_observe(ctx, &ch, cap, (size_t)4 << cap_height));
    CHECK(ctx, bsx_challenger_observe(ctx, &ch, ev, 2 * (size_t)n_cols));
    bsx_batch *oracles[1] = {batch};
    size_t words = 0;
    CHECK(ctx, bsx_fri_proof_words(oracles, 1, &fp, &words));
    uint64_t *proof = malloc(sizeof(uint64_t) * words);
    CHECK(ctx, bsx_fri_prove_openings(ctx, oracles, 1, &inst, &ch, &fp, proof));
    print_words("proof", -1, proof, words);

    /* the prover steps either side of quotient evaluation: Zs + partial products of the permutation argument from
     * the first min(n_cols, 5) columns taken as routed wires (sigma values = a second SplitMix stream), and the
     * quotient hand-over coset_ifft -> chunks -> commitment on two value vectors of 8 n words */
    {
        const int n_routed = n_cols < 5 ? n_cols : 5, max_degree = 2, n_chunks = (n_routed + 1) / 2;
        uint64_t *w = malloc(sizeof(uint64_t) * n * (size_t)n_routed), *sg = malloc(sizeof(uint64_t) * n * (size_t)n_routed);
        for (size_t i = 0; i < n * (size_t)n_routed; i++) {
            w[i] = splitmix(seed, i);
            sg[i] = splitmix(seed + 100, i);
        }
        uint64_t k_is[5], betas[2] = {seed % P + 2, seed % P + 3}, gammas[2] = {seed % P + 4, seed % P + 5};
        for (int j = 0; j < n_routed; j++) k_is[j] = splitmix(seed + 200, (uint64_t)j);
        uint64_t *zs = malloc(sizeof(uint64_t) * n * 2 * (size_t)n_chunks);
        CHECK(ctx, bsx_partial_products_and_zs(ctx, w, sg, BSX_HOST, log_n, n_routed, max_degree, k_is, betas, gammas, 2, zs,
                                               BSX_HOST));
        print_words("zs_row", 1 % (long long)n, NULL, 0);
        for (int c = 0; c < 2 * n_chunks; c++) print_words("zs", c, zs + (size_t)c * n + (1 % n), 1);
        uint64_t zsum = 0;
        for (size_t i = 0; i < n * 2 * (size_t)n_chunks; i++) zsum = zsum * 0x100000001B3ULL ^ zs[i];
        print_words("zs_fnv", -1, &zsum, 1);
        free(w); free(sg); free(zs);

        const int q_bits = 3, factor = 8, q_log_n = log_n > 10 ? 10 : log_n;  /* keeps 2^(log_n + 3) <= 2^24 words small */
        const size_t big = (size_t)1 << (q_log_n + q_bits);
        uint64_t *qv = malloc(sizeof(uint64_t) * 2 * big);
        for (size_t i = 0; i < 2 * big; i++) qv[i] = splitmix(seed + 300, i);
        bsx_batch *qb = NULL;
        CHECK(ctx, bsx_batch_from_quotient_values(ctx, qv, BSX_HOST, q_log_n, q_bits, factor, 2, rate_bits, 0, 0, NULL, &qb, NULL));
        uint64_t qcap[4];
        CHECK(ctx, bsx_batch_cap(qb, qcap));
        print_words("qcap", q_log_n, qcap, 4);
        bsx_batch_free(qb);
        free(qv);
    }

    /* the same batch and proof with the commit sharded over G ranks of one process (all on device 0 here) */
    for (int G = 2; G <= 8; G *= 4) {
        if ((log_n + rate_bits) < 3) break;
        int32_t devs[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        bsx_comm *comm = NULL;
        if (bsx_comm_create(G, devs, &comm) != BSX_OK) {
            fprintf(stderr, "bsx_comm_create failed\n");
            return 1;
        }
        uint64_t *v2 = malloc(sizeof(uint64_t) * n * (size_t)n_cols);
        for (size_t i = 0; i < n * (size_t)n_cols; i++) v2[i] = splitmix(seed, i);
        bsx_sharded_batch *sb = NULL;
        if (bsx_sharded_from_values(comm, v2, log_n, n_cols, rate_bits, cap_height, 0, &sb, NULL) != BSX_OK) {
            fprintf(stderr, "bsx_sharded_from_values: %s\n", bsx_comm_last_error(comm));
            return 1;
        }
        free(v2);
        uint64_t *cap2 = malloc(sizeof(uint64_t) * ((size_t)4 << cap_height));
        bsx_sharded_cap(sb, cap2);
        const int same_cap = memcmp(cap, cap2, sizeof(uint64_t) * ((size_t)4 << cap_height)) == 0;
        int same_rows = bsx_sharded_proof_len(sb) == n_sib;
        uint64_t *row2 = malloc(sizeof(uint64_t) * (size_t)n_cols), *sib2 = malloc(sizeof(uint64_t) * 4 * (size_t)(n_sib + 1));
        for (int k = 0; k < 3; k++) {
            CHECK(ctx, bsx_merkle_get(tree, &leaves[k], 1, row));
            CHECK(ctx, bsx_merkle_prove(tree, leaves[k], sib));
            if (bsx_sharded_get(sb, &leaves[k], 1, row2) != BSX_OK || bsx_sharded_prove(sb, leaves[k], sib2) != BSX_OK) {
                fprintf(stderr, "sharded opening: %s\n", bsx_comm_last_error(comm));
                return 1;
            }
            same_rows = same_rows && memcmp(row, row2, sizeof(uint64_t) * (size_t)n_cols) == 0 &&
                        memcmp(sib, sib2, sizeof(uint64_t) * 4 * (size_t)n_sib) == 0;
        }
        bsx_challenger ch2;
        bsx_challenger_init(&ch2);
        CHECK(ctx, bsx_challenger_observe(ctx, &ch2, cap, (size_t)4 << cap_height));
        CHECK(ctx, bsx_challenger_observe(ctx, &ch2, ev, 2 * (size_t)n_cols));
        uint64_t *proof2 = malloc(sizeof(uint64_t) * words);
        bsx_sharded_batch *sbs[1] = {sb};
        if (bsx_sharded_prove_openings(comm, sbs, 1, &inst, &ch2, &fp, proof2) != BSX_OK) {
            fprintf(stderr, "bsx_sharded_prove_openings: %s\n", bsx_comm_last_error(comm));
            return 1;
        }
        printf("comm %d %d %d %d\n", G, same_cap, same_rows, memcmp(proof, proof2, sizeof(uint64_t) * words) == 0);
        free(proof2); free(row2); free(sib2); free(cap2);
        bsx_sharded_free(sb);
        bsx_comm_destroy(comm);
    }

    /* typed view and bytes of the proof stream */
    {
        bsx_fri_layout lay;
        int32_t width = n_cols;
        if (bsx_fri_proof_layout(&fp, &width, 1, &lay) != BSX_OK) return 1;
        size_t len = 0;
        if (bsx_fri_proof_bytes(proof, &fp, &width, 1, NULL, 0, &len) != BSX_OK) return 1;
        uint8_t *bytes = malloc(len ? len : 1);
        if (bsx_fri_proof_bytes(proof, &fp, &width, 1, bytes, len, &len) != BSX_OK) return 1;
        uint64_t fnv = 0xcbf29ce484222325ULL;
        for (size_t i = 0; i < len; i++) fnv = (fnv ^ bytes[i]) * 0x100000001B3ULL;
        printf("fribytes %zu %016" PRIx64 " %" PRIu64 " %" PRIu64 "\n", len, fnv, lay.total_words, proof[lay.pow_witness_offset]);
        free(bytes);
    }

    /* quotient evaluation with a fourteen-kind gate table (four selector groups) over random committed columns, and two
     * PoseidonGate rows from the device generator */
    {
        const int qn = log_n < 7 ? log_n : 7, nc = 2, routed = 80, wires_w = 135, consts = 6, chunks = 10;
        const size_t rows = (size_t)1 << qn;
        bsx_gate gates[14] = {{BSX_GATE_NOOP, 0, 0, 0, 0, 6}, {BSX_GATE_CONSTANT, 2, 1, 0, 0, 6}, {BSX_GATE_PUBLIC_INPUT, 0, 2, 0, 0, 6},
                              {BSX_GATE_POSEIDON_MDS, 0, 3, 0, 0, 6}, {BSX_GATE_BASE_SUM, 32, 4, 0, 0, 6}, {BSX_GATE_REDUCING, 43, 5, 0, 0, 6},
                              {BSX_GATE_REDUCING_EXTENSION, 32, 6, 1, 6, 11}, {BSX_GATE_ARITHMETIC, 20, 7, 1, 6, 11},
                              {BSX_GATE_ARITHMETIC_EXTENSION, 10, 8, 1, 6, 11}, {BSX_GATE_MUL_EXTENSION, 13, 9, 1, 6, 11},
                              {BSX_GATE_EXPONENTIATION, 66, 10, 1, 6, 11}, {BSX_GATE_RANDOM_ACCESS, 4 | 4 << 8 | 2 << 16, 11, 2, 11, 13},
                              {BSX_GATE_COSET_INTERPOLATION, 4 | 6 << 8, 12, 2, 11, 13}, {BSX_GATE_POSEIDON, 0, 13, 3, 13, 14}};
        uint64_t k_is[80];
        for (int j = 0; j < routed; j++) k_is[j] = splitmix(seed + 400, (uint64_t)j);
        bsx_circuit circ = {qn, 3, 8, nc, wires_w, routed, consts, 4, chunks - 1, 123, 14, gates, k_is};
        const int widths[3] = {consts + routed, wires_w, nc * chunks};
        bsx_batch *b3[3] = {NULL, NULL, NULL};
        for (int k = 0; k < 3; k++) {
            uint64_t *col = malloc(sizeof(uint64_t) * rows * (size_t)widths[k]);
            for (size_t i = 0; i < rows * (size_t)widths[k]; i++) col[i] = splitmix(seed + 500 + (uint64_t)k, i);
            CHECK(ctx, bsx_batch_from_values(ctx, col, BSX_HOST, qn, widths[k], 3, 0, 0, NULL, &b3[k], NULL));
            free(col);
        }
        const uint64_t pih[4] = {seed % P, 1, 2, 3}, betas[2] = {seed % P + 6, seed % P + 7}, gammas[2] = {seed % P + 8, seed % P + 9},
                       alphas[2] = {seed % P + 10, seed % P + 11};
        const size_t qwords = (size_t)nc << (qn + 3);
        uint64_t *qv = malloc(sizeof(uint64_t) * qwords);
        CHECK(ctx, bsx_quotient_values(ctx, &circ, b3[0], b3[1], b3[2], pih, betas, gammas, alphas, qv, BSX_HOST));
        uint64_t fnv = 0;
        for (size_t i = 0; i < qwords; i++) fnv = (fnv * 0x100000001B3ULL) ^ qv[i];
        printf("qv %d %016" PRIx64 " %016" PRIx64 " %016" PRIx64 "\n", qn, fnv, qv[0], qv[qwords - 1]);
        free(qv);
        for (int k = 0; k < 3; k++) bsx_batch_free(b3[k]);
        uint64_t pg[2 * 135];
        for (int i = 0; i < 2 * 135; i++) pg[i] = splitmix(seed + 600, (uint64_t)i);
        pg[24] = 0;
        pg[135 + 24] = 1;
        CHECK(ctx, bsx_poseidon_gate_witness(ctx, pg, BSX_HOST, 2));
        print_words("pgw", -1, pg, 2 * 135);
    }

    /* misuse is reported, not fatal: upstream panics with this message */
    bsx_batch *bad = NULL;
    uint64_t one = 1;
    int32_t rc = bsx_batch_from_values(ctx, &one, BSX_HOST, 0, 1, 1, 5, 0, NULL, &bad, NULL);
    printf("misuse %d %s\n", (int)rc, bsx_last_error(ctx));

    free(proof); free(po); free(pi); free(ev); free(row); free(sib); free(cap);
    bsx_batch_free(batch);
    bsx_ctx_destroy(ctx);
    return 0;
}
