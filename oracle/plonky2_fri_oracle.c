/* plonky2_fri_oracle.c -- CPU restatement of plonky2's Fiat-Shamir challenger and FRI prover /
 * verifier (iop/challenger.rs, fri/oracle.rs prove_openings, fri/prover.rs, fri/verifier.rs).
 * TEST INFRASTRUCTURE ONLY; "parity unpinned" -- see plonky2_oracle.h.  Conventions follow
 * SURVEY.md Appendix A.6/A.7 plus one recalled upstream detail the survey omits: since
 * mir-protocol/plonky2 PR #436 `prove_openings` multiplies the final polynomial by X
 * (`final_poly.coeffs.insert(0, ZERO)`, the quotient is not re-padded) and the verifier multiplies
 * the combined initial evaluation by `subgroup_x`. */
#include "plonky2_oracle.h"
#include <stdlib.h>
#include <string.h>

#define P ORC_P

/* ------------------------------------------------------------ extension helpers */
typedef struct { uint64_t c0, c1; } ext;
static inline ext E(uint64_t a, uint64_t b) { ext r = {a, b}; return r; }
static inline ext eadd(ext a, ext b) { return E(orc_add(a.c0, b.c0), orc_add(a.c1, b.c1)); }
static inline ext esub(ext a, ext b) { return E(orc_sub(a.c0, b.c0), orc_sub(a.c1, b.c1)); }
static inline ext emul(ext a, ext b) {
    uint64_t x[2] = {a.c0, a.c1}, y[2] = {b.c0, b.c1}, o[2];
    orc_ext_mul(x, y, o);
    return E(o[0], o[1]);
}
static inline ext escale(ext a, uint64_t s) { return E(orc_mul(a.c0, s), orc_mul(a.c1, s)); }
static inline ext einv(ext a) {
    uint64_t x[2] = {a.c0, a.c1}, o[2];
    orc_ext_inv(x, o);
    return E(o[0], o[1]);
}
static ext epow(ext a, uint64_t e) {
    ext r = E(1, 0);
    while (e) {
        if (e & 1) r = emul(r, a);
        a = emul(a, a);
        e >>= 1;
    }
    return r;
}
static inline int eeq(ext a, ext b) { return a.c0 == b.c0 && a.c1 == b.c1; }

/* ------------------------------------------------------------ challenger (iop/challenger.rs) */
void orc_challenger_init(orc_challenger *c) { memset(c, 0, sizeof(*c)); }
static void duplexing(orc_challenger *c) {
    /* overwrite mode: the buffered inputs replace the first lanes of the sponge state */
    for (int i = 0; i < c->n_in; i++) c->state[i] = c->in[i];
    c->n_in = 0;
    orc_poseidon(c->state);
    memcpy(c->out, c->state, 8 * sizeof(uint64_t));
    c->n_out = 8;
}
void orc_challenger_observe(orc_challenger *c, uint64_t x) {
    c->n_out = 0; /* buffered outputs are now invalid */
    c->in[c->n_in++] = x;
    if (c->n_in == 8) duplexing(c);
}
void orc_challenger_observe_many(orc_challenger *c, const uint64_t *x, size_t n) {
    for (size_t i = 0; i < n; i++) orc_challenger_observe(c, x[i]);
}
uint64_t orc_challenger_get(orc_challenger *c) {
    if (c->n_in != 0 || c->n_out == 0) duplexing(c);
    return c->out[--c->n_out]; /* Vec::pop: from the back */
}
static ext challenger_get_ext(orc_challenger *c) {
    uint64_t a = orc_challenger_get(c), b = orc_challenger_get(c);
    return E(a, b);
}
void orc_challenger_get_ext(orc_challenger *c, uint64_t out[2]) {
    ext e = challenger_get_ext(c);
    out[0] = e.c0;
    out[1] = e.c1;
}

/* ------------------------------------------------------------ extension FFT helpers */
/* component-wise coset FFT of an extension polynomial stored interleaved [c0, c1, ...] */
static void ext_coset_fft(uint64_t *a, int log_n, uint64_t shift) {
    size_t n = (size_t)1 << log_n;
    uint64_t s = 1;
    for (size_t i = 0; i < n; i++) {
        a[2 * i] = orc_mul(a[2 * i], s);
        a[2 * i + 1] = orc_mul(a[2 * i + 1], s);
        s = orc_mul(s, shift);
    }
    orc_fft_strided(a, log_n, 2);
    orc_fft_strided(a + 1, log_n, 2);
}

/* ------------------------------------------------------------ prove_openings: final polynomial
 * fri/oracle.rs PolynomialBatch::prove_openings up to (not including) the LDE.  out: 2^log_n
 * extension coefficients. */
int orc_fri_final_poly(const orc_batch_view *oracles, int n_oracles, const orc_fri_instance *inst,
                       const uint64_t alpha_in[2], uint64_t *out) {
    if (n_oracles <= 0) return -1;
    int log_n = oracles[0].log_n;
    size_t n = (size_t)1 << log_n;
    ext alpha = E(alpha_in[0], alpha_in[1]);
    ext *fin = (ext *)calloc(n, sizeof(ext));
    ext *comp = (ext *)malloc(n * sizeof(ext));
    int pos = 0;
    for (int b = 0; b < inst->n_batches; b++) {
        ext z = E(inst->points[2 * b], inst->points[2 * b + 1]);
        /* ReducingFactor::reduce_polys_base: sum_j alpha^j * p_j, count = number of polys */
        for (size_t i = 0; i < n; i++) comp[i] = E(0, 0);
        ext ap = E(1, 0);
        int cnt = inst->n_polys[b];
        for (int j = 0; j < cnt; j++, pos++) {
            const orc_batch_view *o = &oracles[inst->poly_oracle[pos]];
            if (o->log_n != log_n) { free(fin); free(comp); return -1; }
            const uint64_t *p = o->coeffs + (size_t)inst->poly_index[pos] * n;
            for (size_t i = 0; i < n; i++) comp[i] = eadd(comp[i], escale(ap, p[i]));
            ap = emul(ap, alpha);
        }
        /* alpha.shift_poly(&mut final_poly): final_poly *= alpha^count */
        ext sh = epow(alpha, (uint64_t)cnt);
        for (size_t i = 0; i < n; i++) fin[i] = emul(fin[i], sh);
        /* divide_by_linear(z): quotient q_{i-1} = acc_i, acc_i = acc_{i+1} * z + c_i (from the top);
         * final_poly += quotient (length n-1) */
        ext acc = E(0, 0);
        for (size_t i = n; i-- > 0;) {
            acc = eadd(emul(acc, z), comp[i]);
            if (i >= 1) fin[i - 1] = eadd(fin[i - 1], acc);
        }
    }
    /* final_poly.coeffs.insert(0, ZERO): multiply by X (PR #436); degree n-2 quotient -> n coefficients */
    out[0] = out[1] = 0;
    for (size_t i = 0; i + 1 < n; i++) {
        out[2 * (i + 1)] = fin[i].c0;
        out[2 * (i + 1) + 1] = fin[i].c1;
    }
    free(fin);
    free(comp);
    return 0;
}

/* ------------------------------------------------------------ proof layout
 * flat u64 stream in plonky2 FriProof field order:
 *   commit_phase_merkle_caps[n_rounds][2^cap][4]
 *   query_round_proofs[num_query_rounds]:
 *       initial_trees_proof: per oracle: leaf[width], siblings[log_nb - cap][4]
 *       steps: per round i: evals[2^arity_i][2], siblings[log_leaves_i - cap][4]
 *   final_poly[final_len][2]
 *   pow_witness */
size_t orc_fri_proof_words(const orc_batch_view *oracles, int n_oracles, const orc_fri_params *fp) {
    int log_nb = fp->degree_bits + fp->rate_bits, cap = fp->cap_height;
    size_t w = 0;
    w += (size_t)fp->n_arity * ((size_t)4 << cap);
    size_t per_q = 0;
    for (int o = 0; o < n_oracles; o++)
        per_q += (size_t)(oracles[o].n_cols + oracles[o].n_salt) + 4 * (size_t)(log_nb - cap);
    int lg = log_nb;
    for (int i = 0; i < fp->n_arity; i++) {
        lg -= fp->arity_bits[i];
        per_q += ((size_t)2 << fp->arity_bits[i]) + 4 * (size_t)(lg - cap);
    }
    w += per_q * (size_t)fp->num_query_rounds;
    int tot = 0;
    for (int i = 0; i < fp->n_arity; i++) tot += fp->arity_bits[i];
    w += (size_t)2 << (fp->degree_bits - tot);
    return w + 1;
}

/* ------------------------------------------------------------ fri_proof (fri/prover.rs) */
static void observe_cap(orc_challenger *ch, const uint64_t *cap, int cap_height) {
    orc_challenger_observe_many(ch, cap, (size_t)4 << cap_height);
}

int orc_fri_prove_openings(const orc_batch_view *oracles, int n_oracles, const orc_fri_instance *inst,
                           orc_challenger *ch, const orc_fri_params *fp, uint64_t *proof) {
    const int log_n = fp->degree_bits, rb = fp->rate_bits, log_nb = log_n + rb, cap = fp->cap_height;
    const size_t n = (size_t)1 << log_n, nb = (size_t)1 << log_nb;
    uint64_t alpha[2];
    orc_challenger_get_ext(ch, alpha);
    /* coefficients of the LDE polynomial: first 1/rate are non-zero */
    uint64_t *coeffs = (uint64_t *)calloc(2 * nb, sizeof(uint64_t));
    uint64_t *values = (uint64_t *)malloc(2 * nb * sizeof(uint64_t));
    if (orc_fri_final_poly(oracles, n_oracles, inst, alpha, coeffs) != 0) { free(coeffs); free(values); return -1; }
    (void)n;
    memcpy(values, coeffs, 2 * nb * sizeof(uint64_t));
    ext_coset_fft(values, log_nb, 7);

    /* ---- fri_committed_trees */
    int n_rounds = fp->n_arity;
    uint64_t **tree_leaves = (uint64_t **)calloc(n_rounds ? n_rounds : 1, sizeof(uint64_t *));
    uint64_t **tree_digests = (uint64_t **)calloc(n_rounds ? n_rounds : 1, sizeof(uint64_t *));
    int *tree_log = (int *)calloc(n_rounds ? n_rounds : 1, sizeof(int));
    uint64_t *w = proof;
    size_t len = nb;
    int lg = log_nb;
    uint64_t shift = 7;
    for (int r = 0; r < n_rounds; r++) {
        int ab = fp->arity_bits[r];
        size_t arity = (size_t)1 << ab;
        /* reverse_index_bits_in_place(values); leaves = chunks of `arity` extension values, flattened */
        uint64_t *lv = (uint64_t *)malloc(2 * len * sizeof(uint64_t));
        for (size_t i = 0; i < len; i++) {
            size_t j = orc_reverse_bits(i, lg);
            lv[2 * i] = values[2 * j];
            lv[2 * i + 1] = values[2 * j + 1];
        }
        int tl = lg - ab;
        size_t nd = 4 * 2 * (((size_t)1 << tl) - ((size_t)1 << cap));
        uint64_t *dg = (uint64_t *)malloc((nd ? nd : 1) * sizeof(uint64_t));
        if (orc_merkle_build(lv, tl, 2 * arity, cap, nd ? dg : NULL, w) != 0) return -1;
        observe_cap(ch, w, cap);
        tree_leaves[r] = lv;
        tree_digests[r] = dg;
        tree_log[r] = tl;
        w += (size_t)4 << cap;
        ext beta = challenger_get_ext(ch);
        /* coeffs <- chunks_exact(arity).map(reduce_with_powers(chunk, beta)) */
        size_t nl = len >> ab;
        for (size_t j = 0; j < nl; j++) {
            ext acc = E(0, 0);
            for (size_t k = arity; k-- > 0;)
                acc = eadd(emul(acc, beta), E(coeffs[2 * (arity * j + k)], coeffs[2 * (arity * j + k) + 1]));
            coeffs[2 * j] = acc.c0;
            coeffs[2 * j + 1] = acc.c1;
        }
        shift = orc_pow(shift, arity);
        len = nl;
        lg -= ab;
        memcpy(values, coeffs, 2 * len * sizeof(uint64_t));
        ext_coset_fft(values, lg, shift);
    }
    /* coeffs.truncate(len >> rate_bits); challenger.observe_extension_elements(coeffs) */
    size_t final_len = len >> rb;
    orc_challenger_observe_many(ch, coeffs, 2 * final_len);

    /* ---- fri_proof_of_work: smallest witness whose response has >= pow_bits leading zeros */
    uint64_t pow_witness = 0;
    for (uint64_t cand = 0;; cand++) {
        orc_challenger c2 = *ch;
        orc_challenger_observe(&c2, cand);
        uint64_t resp = orc_challenger_get(&c2);
        int lz = resp == 0 ? 64 : __builtin_clzll(resp);
        if (lz >= fp->proof_of_work_bits) {
            pow_witness = cand;
            break;
        }
    }
    orc_challenger_observe(ch, pow_witness);
    (void)orc_challenger_get(ch); /* pow_response */

    /* ---- fri_prover_query_rounds */
    for (int q = 0; q < fp->num_query_rounds; q++) {
        uint64_t rnd = orc_challenger_get(ch);
        size_t x = (size_t)(rnd % nb);
        for (int o = 0; o < n_oracles; o++) {
            size_t width = (size_t)(oracles[o].n_cols + oracles[o].n_salt);
            memcpy(w, oracles[o].leaves + x * width, width * sizeof(uint64_t));
            w += width;
            if (orc_merkle_prove(oracles[o].digests, log_nb, cap, x, w) != 0) return -1;
            w += 4 * (size_t)(log_nb - cap);
        }
        for (int r = 0; r < n_rounds; r++) {
            int ab = fp->arity_bits[r];
            size_t li = x >> ab, lw = (size_t)2 << ab;
            memcpy(w, tree_leaves[r] + li * lw, lw * sizeof(uint64_t));
            w += lw;
            if (orc_merkle_prove(tree_digests[r], tree_log[r], cap, li, w) != 0) return -1;
            w += 4 * (size_t)(tree_log[r] - cap);
            x = li;
        }
    }
    memcpy(w, coeffs, 2 * final_len * sizeof(uint64_t));
    w += 2 * final_len;
    *w++ = pow_witness;
    for (int r = 0; r < n_rounds; r++) {
        free(tree_leaves[r]);
        free(tree_digests[r]);
    }
    free(tree_leaves);
    free(tree_digests);
    free(tree_log);
    free(coeffs);
    free(values);
    return 0;
}

/* ------------------------------------------------------------ verifier (fri/verifier.rs) */
void orc_eval_poly_ext(const orc_batch_view *o, int poly_index, const uint64_t z[2], uint64_t out[2]) {
    size_t n = (size_t)1 << o->log_n;
    const uint64_t *p = o->coeffs + (size_t)poly_index * n;
    ext acc = E(0, 0), zz = E(z[0], z[1]);
    for (size_t i = n; i-- > 0;) acc = eadd(emul(acc, zz), E(p[i], 0));
    out[0] = acc.c0;
    out[1] = acc.c1;
}

/* value at `at` of the polynomial of degree < k through (xs[t], ys[t]) -- plain Lagrange form
 * (upstream uses barycentric weights in fri/verifier.rs compute_evaluation; same polynomial) */
static ext interpolate_at(const ext *xs, const ext *ys, int k, ext at) {
    ext sum = E(0, 0);
    for (int i = 0; i < k; i++) {
        ext num = E(1, 0), den = E(1, 0);
        for (int j = 0; j < k; j++)
            if (j != i) {
                num = emul(num, esub(at, xs[j]));
                den = emul(den, esub(xs[i], xs[j]));
            }
        sum = eadd(sum, emul(ys[i], emul(num, einv(den))));
    }
    return sum;
}

int orc_fri_verify(const orc_batch_view *oracles, int n_oracles, const orc_fri_instance *inst,
                   const uint64_t *openings, orc_challenger *ch, const orc_fri_params *fp,
                   const uint64_t *proof) {
    const int log_n = fp->degree_bits, rb = fp->rate_bits, log_nb = log_n + rb, cap = fp->cap_height;
    const size_t nb = (size_t)1 << log_nb;
    const int n_rounds = fp->n_arity;
    /* ---- locate the proof sections */
    const uint64_t *caps = proof;
    const uint64_t *queries = caps + (size_t)n_rounds * ((size_t)4 << cap);
    size_t per_q = 0;
    for (int o = 0; o < n_oracles; o++)
        per_q += (size_t)(oracles[o].n_cols + oracles[o].n_salt) + 4 * (size_t)(log_nb - cap);
    int lg = log_nb, tot = 0;
    for (int i = 0; i < n_rounds; i++) {
        lg -= fp->arity_bits[i];
        tot += fp->arity_bits[i];
        per_q += ((size_t)2 << fp->arity_bits[i]) + 4 * (size_t)(lg - cap);
    }
    const uint64_t *final_poly = queries + per_q * (size_t)fp->num_query_rounds;
    size_t final_len = (size_t)1 << (log_n - tot);
    uint64_t pow_witness = final_poly[2 * final_len];

    /* ---- Challenger::fri_challenges */
    ext alpha = challenger_get_ext(ch);
    ext betas[16];
    for (int r = 0; r < n_rounds; r++) {
        observe_cap(ch, caps + (size_t)r * ((size_t)4 << cap), cap);
        betas[r] = challenger_get_ext(ch);
    }
    orc_challenger_observe_many(ch, final_poly, 2 * final_len);
    orc_challenger_observe(ch, pow_witness);
    uint64_t pow_response = orc_challenger_get(ch);
    int lz = pow_response == 0 ? 64 : __builtin_clzll(pow_response);
    if (lz < fp->proof_of_work_bits) return 0;

    /* ---- PrecomputedReducedOpenings::from_os_and_alpha */
    ext reduced_openings[16];
    int pos = 0;
    for (int b = 0; b < inst->n_batches; b++) {
        ext acc = E(0, 0), ap = E(1, 0);
        for (int j = 0; j < inst->n_polys[b]; j++, pos++) {
            acc = eadd(acc, emul(ap, E(openings[2 * pos], openings[2 * pos + 1])));
            ap = emul(ap, alpha);
        }
        reduced_openings[b] = acc;
    }

    for (int q = 0; q < fp->num_query_rounds; q++) {
        size_t x_index = (size_t)(orc_challenger_get(ch) % nb);
        const uint64_t *w = queries + per_q * (size_t)q;
        /* fri_verify_initial_proof */
        const uint64_t *rows[64];
        for (int o = 0; o < n_oracles; o++) {
            size_t width = (size_t)(oracles[o].n_cols + oracles[o].n_salt);
            rows[o] = w;
            if (!orc_merkle_verify(w, width, x_index, w + width, log_nb - cap, oracles[o].cap, cap)) return 0;
            w += width + 4 * (size_t)(log_nb - cap);
        }
        uint64_t sx = orc_mul(7, orc_pow(orc_root_of_unity(log_nb), orc_reverse_bits(x_index, log_nb)));
        ext subgroup_x = E(sx, 0);
        /* fri_combine_initial */
        ext sum = E(0, 0);
        pos = 0;
        for (int b = 0; b < inst->n_batches; b++) {
            ext acc = E(0, 0), ap = E(1, 0);
            int cnt = inst->n_polys[b];
            for (int j = 0; j < cnt; j++, pos++) {
                uint64_t e = rows[inst->poly_oracle[pos]][inst->poly_index[pos]]; /* unsalted_eval */
                acc = eadd(acc, escale(ap, e));
                ap = emul(ap, alpha);
            }
            ext num = esub(acc, reduced_openings[b]);
            ext den = esub(subgroup_x, E(inst->points[2 * b], inst->points[2 * b + 1]));
            sum = emul(sum, epow(alpha, (uint64_t)cnt));
            sum = eadd(sum, emul(num, einv(den)));
        }
        ext old_eval = emul(sum, subgroup_x); /* final polynomial was multiplied by X */
        /* reduction steps */
        int cur_log = log_nb;
        uint64_t shift = 7;
        for (int r = 0; r < n_rounds; r++) {
            int ab = fp->arity_bits[r], arity = 1 << ab;
            const uint64_t *evals = w;
            size_t coset_index = x_index >> ab, within = x_index & (size_t)(arity - 1);
            if (!eeq(E(evals[2 * within], evals[2 * within + 1]), old_eval)) return 0;
            /* the leaf holds P_r at shift_r * omega^{bitrev(arity * coset_index + t)} */
            ext xs[64], ys[64];
            uint64_t wr = orc_root_of_unity(cur_log);
            for (int t = 0; t < arity; t++) {
                uint64_t e = orc_reverse_bits((uint64_t)arity * coset_index + (uint64_t)t, cur_log);
                xs[t] = E(orc_mul(shift, orc_pow(wr, e)), 0);
                ys[t] = E(evals[2 * t], evals[2 * t + 1]);
            }
            old_eval = interpolate_at(xs, ys, arity, betas[r]);
            int tl = cur_log - ab;
            if (!orc_merkle_verify(evals, (size_t)2 * arity, coset_index, evals + 2 * arity, tl - cap,
                                   caps + (size_t)r * ((size_t)4 << cap), cap))
                return 0;
            w += (size_t)2 * arity + 4 * (size_t)(tl - cap);
            subgroup_x = epow(subgroup_x, (uint64_t)arity);
            shift = orc_pow(shift, (uint64_t)arity);
            x_index = coset_index;
            cur_log = tl;
        }
        /* final_poly.eval(subgroup_x) == old_eval */
        ext acc = E(0, 0);
        for (size_t i = final_len; i-- > 0;)
            acc = eadd(emul(acc, subgroup_x), E(final_poly[2 * i], final_poly[2 * i + 1]));
        if (!eeq(acc, old_eval)) return 0;
    }
    return 1;
}
