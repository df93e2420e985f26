/* plonky2_oracle.c -- CPU restatement of the plonky2 commitment phase.
 * TEST INFRASTRUCTURE ONLY; "parity unpinned" against the real prover -- see plonky2_oracle.h.
 * Upstream file names are given per function; the reference mount is empty so no
 * file:line into /root/reference exists. */
#include "plonky2_oracle.h"
#include "poseidon_rc.h"
#include <stdlib.h>
#include <string.h>
#ifdef __AVX2__
#include <immintrin.h>
#endif
#ifdef _OPENMP
#include <omp.h>
#endif

typedef unsigned __int128 u128;
#define P ORC_P
#define EPS 0xFFFFFFFFULL

/* ------------------------------------------------------------------ field */
/* plonky2 field/src/goldilocks_field.rs: Add/Sub/Mul, reduce128; SURVEY Appendix A.1 */
static inline uint64_t canon(uint64_t x) {
    uint64_t y = x - P;
    return x >= P ? y : x;
}
/* branch-free: data-dependent branches mispredict on random field elements */
static inline uint64_t add_(uint64_t a, uint64_t b) {
    uint64_t s;
    uint64_t cy = __builtin_add_overflow(a, b, &s);
    s += EPS & (0 - cy); /* wrapped 2^64 == EPS (mod p) */
    return canon(s);
}
static inline uint64_t sub_(uint64_t a, uint64_t b) {
    uint64_t d;
    uint64_t br = __builtin_sub_overflow(a, b, &d);
    return d - (EPS & (0 - br)); /* borrowed 2^64 == EPS */
}
static inline uint64_t reduce128(u128 x) {
    uint64_t lo = (uint64_t)x, hi = (uint64_t)(x >> 64);
    uint64_t hh = hi >> 32, hl = hi & EPS;
    uint64_t t0, r;
    uint64_t br = __builtin_sub_overflow(lo, hh, &t0);
    t0 -= EPS & (0 - br);  /* borrow: -2^64 == -EPS */
    uint64_t t1 = hl * EPS; /* 2^64*hl == EPS*hl */
    uint64_t cy = __builtin_add_overflow(t0, t1, &r);
    r += EPS & (0 - cy);
    return canon(r);
}
static inline uint64_t mul_(uint64_t a, uint64_t b) { return reduce128((u128)a * b); }

uint64_t orc_add(uint64_t a, uint64_t b) { return add_(canon(a), canon(b)); }
uint64_t orc_sub(uint64_t a, uint64_t b) { return sub_(canon(a), canon(b)); }
uint64_t orc_mul(uint64_t a, uint64_t b) { return mul_(a, b); }
uint64_t orc_pow(uint64_t a, uint64_t e) {
    uint64_t r = 1;
    while (e) {
        if (e & 1) r = mul_(r, a);
        a = mul_(a, a);
        e >>= 1;
    }
    return r;
}
uint64_t orc_inv(uint64_t a) { return orc_pow(a, P - 2); }
/* POWER_OF_TWO_GENERATOR = 7^((p-1)/2^32); primitive_root_of_unity(n) = that ^ 2^(32-n) */
uint64_t orc_root_of_unity(int log_n) {
    uint64_t w = 1753635133440165772ULL;
    for (int i = log_n; i < 32; i++) w = mul_(w, w);
    return w;
}
/* field/src/extension/quadratic.rs: W = 7 */
void orc_ext_mul(const uint64_t a[2], const uint64_t b[2], uint64_t out[2]) {
    uint64_t c0 = add_(mul_(a[0], b[0]), mul_(7, mul_(a[1], b[1])));
    uint64_t c1 = add_(mul_(a[0], b[1]), mul_(a[1], b[0]));
    out[0] = c0;
    out[1] = c1;
}
void orc_ext_inv(const uint64_t a[2], uint64_t out[2]) {
    /* 1/(a0 + a1 x) = (a0 - a1 x) / (a0^2 - 7 a1^2) */
    uint64_t n = sub_(mul_(a[0], a[0]), mul_(7, mul_(a[1], a[1])));
    uint64_t ni = orc_inv(n);
    out[0] = mul_(a[0], ni);
    out[1] = mul_(sub_(0, a[1]), ni);
}

/* -------------------------------------------------------------------- FFT */
uint64_t orc_reverse_bits(uint64_t x, int bits) {
    uint64_t r = 0;
    for (int i = 0; i < bits; i++) r |= ((x >> i) & 1ULL) << (bits - 1 - i);
    return r;
}

/* field/src/fft.rs fft_dispatch: reverse_index_bits_in_place, then classic DIT
 * butterflies with omega = primitive_root_of_unity(log_n).  Natural in, natural out. */
static void fft_core(uint64_t *a, int log_n, size_t stride, uint64_t omega) {
    size_t n = (size_t)1 << log_n;
    for (size_t i = 0; i < n; i++) {
        size_t j = orc_reverse_bits(i, log_n);
        if (i < j) {
            uint64_t t = a[i * stride];
            a[i * stride] = a[j * stride];
            a[j * stride] = t;
        }
    }
    if (log_n == 0) return;
    uint64_t *tw = (uint64_t *)malloc(sizeof(uint64_t) * (n / 2 ? n / 2 : 1));
    tw[0] = 1;
    for (size_t i = 1; i < n / 2; i++) tw[i] = mul_(tw[i - 1], omega);
    for (int s = 1; s <= log_n; s++) {
        size_t half = (size_t)1 << (s - 1), step = n >> s;
        for (size_t k = 0; k < n; k += 2 * half)
            for (size_t j = 0; j < half; j++) {
                uint64_t u = a[(k + j) * stride];
                uint64_t v = mul_(a[(k + j + half) * stride], tw[j * step]);
                a[(k + j) * stride] = add_(u, v);
                a[(k + j + half) * stride] = sub_(u, v);
            }
    }
    free(tw);
}
void orc_fft_strided(uint64_t *a, int log_n, size_t stride) {
    fft_core(a, log_n, stride, orc_root_of_unity(log_n));
}
void orc_fft(uint64_t *a, int log_n) { fft_core(a, log_n, 1, orc_root_of_unity(log_n)); }
/* field/src/fft.rs ifft_with_options: forward transform with omega^-1, scaled by n^-1 */
void orc_ifft(uint64_t *a, int log_n) {
    size_t n = (size_t)1 << log_n;
    fft_core(a, log_n, 1, orc_inv(orc_root_of_unity(log_n)));
    uint64_t ninv = orc_inv((uint64_t)n % P);
    for (size_t i = 0; i < n; i++) a[i] = mul_(a[i], ninv);
}
/* field/src/polynomial/mod.rs coset_fft_with_options: c_j *= shift^j, then fft */
void orc_coset_fft(uint64_t *a, int log_n, uint64_t shift) {
    size_t n = (size_t)1 << log_n;
    uint64_t s = 1;
    for (size_t i = 0; i < n; i++) {
        a[i] = mul_(a[i], s);
        s = mul_(s, shift);
    }
    orc_fft(a, log_n);
}
/* field/src/polynomial/mod.rs PolynomialValues::coset_ifft: ifft, then c_j *= shift^-j */
void orc_coset_ifft(uint64_t *a, int log_n, uint64_t shift) {
    size_t n = (size_t)1 << log_n;
    orc_ifft(a, log_n);
    uint64_t sinv = orc_inv(shift % P), s = 1;
    for (size_t i = 0; i < n; i++) {
        a[i] = mul_(a[i], s);
        s = mul_(s, sinv);
    }
}
/* PolynomialCoeffs::lde (zero pad) followed by coset_fft (fri/oracle.rs lde_values) */
void orc_lde(const uint64_t *coeffs, int log_n, int rate_bits, uint64_t shift, uint64_t *out) {
    size_t n = (size_t)1 << log_n, nb = n << rate_bits;
    memcpy(out, coeffs, n * sizeof(uint64_t));
    memset(out + n, 0, (nb - n) * sizeof(uint64_t));
    orc_coset_fft(out, log_n + rate_bits, shift);
}

/* --------------------------------------------------------------- Poseidon */
/* hash/poseidon.rs + poseidon_goldilocks.rs; naive-round form, SURVEY Appendix A.3 */
static const uint64_t RC[POSEIDON_N_RC] = {POSEIDON_RC_LIST};
static const uint64_t MDS_CIRC[12] = {17, 15, 41, 16, 2, 28, 13, 13, 39, 18, 34, 20};
static const uint64_t MDS_DIAG0 = 8;
const uint64_t *orc_round_constants(void) { return RC; }

static inline uint64_t sbox(uint64_t x) {
    uint64_t x2 = mul_(x, x), x4 = mul_(x2, x2), x3 = mul_(x, x2);
    return mul_(x3, x4);
}
static inline void mds_layer(uint64_t s[12]) {
    /* Split each word into 32-bit halves so the 12-term dot products with the small
     * circulant constants stay inside u64 (12 * 41 * 2^32 < 2^42); recombine and reduce once.
     * Same value as the direct sum  sum_i CIRC[i]*s[(i+r)%12] + DIAG[r]*s[r]  (Appendix A.3). */
    uint64_t lo[24], hi[24];
    uint64_t al[12], ah[12];
    for (int i = 0; i < 12; i++) {
        lo[i] = lo[i + 12] = s[i] & EPS;
        hi[i] = hi[i + 12] = s[i] >> 32;
    }
#ifdef __AVX2__
    for (int b = 0; b < 12; b += 4) {
        __m256i vl = _mm256_setzero_si256(), vh = _mm256_setzero_si256();
        for (int i = 0; i < 12; i++) {
            __m256i c = _mm256_set1_epi64x((long long)MDS_CIRC[i]);
            vl = _mm256_add_epi64(vl, _mm256_mul_epu32(c, _mm256_loadu_si256((const __m256i *)(lo + i + b))));
            vh = _mm256_add_epi64(vh, _mm256_mul_epu32(c, _mm256_loadu_si256((const __m256i *)(hi + i + b))));
        }
        _mm256_storeu_si256((__m256i *)(al + b), vl);
        _mm256_storeu_si256((__m256i *)(ah + b), vh);
    }
#else
    for (int r = 0; r < 12; r++) {
        al[r] = ah[r] = 0;
        for (int i = 0; i < 12; i++) {
            al[r] += MDS_CIRC[i] * lo[i + r];
            ah[r] += MDS_CIRC[i] * hi[i + r];
        }
    }
#endif
    al[0] += MDS_DIAG0 * lo[0];
    ah[0] += MDS_DIAG0 * hi[0];
    for (int r = 0; r < 12; r++) s[r] = reduce128((u128)al[r] + ((u128)ah[r] << 32));
}
uint64_t orc_poseidon_sbox(uint64_t x) { return sbox(canon(x)); }
void orc_poseidon_mds(uint64_t s[12]) { mds_layer(s); }
void orc_poseidon(uint64_t s[12]) {
    for (int i = 0; i < 12; i++) s[i] = canon(s[i]);
    for (int r = 0; r < 30; r++) {
        for (int i = 0; i < 12; i++) s[i] = add_(s[i], RC[12 * r + i]);
        if (r < 4 || r >= 26)
            for (int i = 0; i < 12; i++) s[i] = sbox(s[i]);
        else
            s[0] = sbox(s[0]);
        mds_layer(s);
    }
}
/* hash/hashing.rs hash_n_to_m_no_pad: overwrite-mode sponge, rate 8, no padding */
void orc_hash_no_pad(const uint64_t *in, size_t len, uint64_t out[4]) {
    uint64_t st[12] = {0};
    for (size_t off = 0; off < len; off += 8) {
        size_t k = len - off < 8 ? len - off : 8;
        for (size_t i = 0; i < k; i++) st[i] = in[off + i];
        orc_poseidon(st);
    }
    memcpy(out, st, 4 * sizeof(uint64_t));
}
/* hash/hashing.rs hash_or_noop: <= 4 elements are zero-extended, not hashed */
void orc_hash_or_noop(const uint64_t *in, size_t len, uint64_t out[4]) {
    if (len <= 4) {
        for (size_t i = 0; i < 4; i++) out[i] = i < len ? in[i] : 0;
    } else
        orc_hash_no_pad(in, len, out);
}
/* hash/hashing.rs compress / PoseidonHash::two_to_one */
void orc_two_to_one(const uint64_t l[4], const uint64_t r[4], uint64_t out[4]) {
    uint64_t st[12] = {l[0], l[1], l[2], l[3], r[0], r[1], r[2], r[3], 0, 0, 0, 0};
    orc_poseidon(st);
    memcpy(out, st, 4 * sizeof(uint64_t));
}

/* ----------------------------------------------------------------- Merkle */
/* hash/merkle_tree.rs MerkleTree::new: leaf digests by hash_or_noop, parents by
 * two_to_one, down to 2^cap_height roots.  Level-order storage (see header). */
int orc_merkle_build(const uint64_t *leaves, int log_leaves, size_t leaf_len, int cap_height,
                     uint64_t *digests, uint64_t *cap) {
    if (log_leaves < 0 || cap_height < 0 || cap_height > log_leaves) return -1;
    size_t n = (size_t)1 << log_leaves;
    int levels = log_leaves - cap_height; /* stored levels below the cap */
    uint64_t *tmp = NULL;
    if (!digests && levels > 0) {
        tmp = (uint64_t *)malloc(sizeof(uint64_t) * 4 * 2 * (n - ((size_t)1 << cap_height)));
        digests = tmp;
    }
    uint64_t *lvl = levels > 0 ? digests : cap;
#pragma omp parallel for schedule(static)
    for (size_t i = 0; i < n; i++) orc_hash_or_noop(leaves + i * leaf_len, leaf_len, lvl + 4 * i);
    size_t cnt = n;
    for (int k = 1; k <= levels; k++) {
        uint64_t *nxt = k == levels ? cap : lvl + 4 * cnt;
        size_t half = cnt / 2;
#pragma omp parallel for schedule(static) if (half > 256)
        for (size_t i = 0; i < half; i++) orc_two_to_one(lvl + 8 * i, lvl + 8 * i + 4, nxt + 4 * i);
        lvl = nxt;
        cnt = half;
    }
    free(tmp);
    return 0;
}
/* hash/merkle_tree.rs MerkleTree::prove: sibling digests bottom-up, below the cap */
int orc_merkle_prove(const uint64_t *digests, int log_leaves, int cap_height, size_t leaf_index,
                     uint64_t *siblings) {
    if (cap_height > log_leaves || leaf_index >> log_leaves) return -1;
    size_t n = (size_t)1 << log_leaves, off = 0, idx = leaf_index;
    for (int k = 0; k < log_leaves - cap_height; k++) {
        memcpy(siblings + 4 * k, digests + 4 * (off + (idx ^ 1)), 4 * sizeof(uint64_t));
        off += n >> k;
        idx >>= 1;
    }
    return 0;
}
/* hash/merkle_proofs.rs verify_merkle_proof_to_cap */
int orc_merkle_verify(const uint64_t *leaf, size_t leaf_len, size_t leaf_index,
                      const uint64_t *siblings, int n_siblings, const uint64_t *cap, int cap_height) {
    uint64_t cur[4], nxt[4];
    orc_hash_or_noop(leaf, leaf_len, cur);
    size_t idx = leaf_index;
    for (int k = 0; k < n_siblings; k++) {
        if (idx & 1)
            orc_two_to_one(siblings + 4 * k, cur, nxt);
        else
            orc_two_to_one(cur, siblings + 4 * k, nxt);
        memcpy(cur, nxt, sizeof(cur));
        idx >>= 1;
    }
    if (idx >> cap_height) return 0;
    return memcmp(cur, cap + 4 * idx, sizeof(cur)) == 0;
}
/* hash/merkle_tree.rs fill_subtree layout: per cap subtree, a buffer of 2^(k+1)-2 digests
 * = left recursive buffer || left child digest || right child digest || right recursive buffer. */
static void fill_p2(const uint64_t *lv, int log_leaves, size_t n, int height, size_t node,
                    uint64_t *buf, size_t buf_len) {
    /* subtree rooted at (level `height`, index `node`); buf receives its 2^(height+1)-2 digests */
    if (buf_len == 0) return;
    size_t half = buf_len / 2, off = 0;
    for (int k = 0; k < height - 1; k++) off += n >> k; /* offset of level height-1 */
    (void)log_leaves;
    memcpy(buf + 4 * (half - 1), lv + 4 * (off + 2 * node), 4 * sizeof(uint64_t));
    memcpy(buf + 4 * half, lv + 4 * (off + 2 * node + 1), 4 * sizeof(uint64_t));
    fill_p2(lv, log_leaves, n, height - 1, 2 * node, buf, half - 1);
    fill_p2(lv, log_leaves, n, height - 1, 2 * node + 1, buf + 4 * (half + 1), half - 1);
}
void orc_merkle_digests_to_plonky2(const uint64_t *level_digests, int log_leaves, int cap_height,
                                   uint64_t *out) {
    size_t n = (size_t)1 << log_leaves;
    int height = log_leaves - cap_height;
    size_t sub = ((size_t)2 << height) - 2;
    for (size_t c = 0; c < ((size_t)1 << cap_height); c++)
        fill_p2(level_digests, log_leaves, n, height, c, out + 4 * c * sub, sub);
}

/* -------------------------------------------------------- PolynomialBatch */
/* fri/oracle.rs from_coeffs: lde_values -> transpose -> reverse_index_bits_in_place -> MerkleTree::new */
int orc_batch_from_coeffs(const uint64_t *coeffs, int log_n, int n_cols, int rate_bits,
                          int cap_height, const uint64_t *salt, int n_salt, uint64_t *leaves,
                          uint64_t *digests, uint64_t *cap) {
    int log_nb = log_n + rate_bits;
    if (log_n < 0 || n_cols < 0 || rate_bits < 0 || cap_height > log_nb) return -1;
    size_t n = (size_t)1 << log_n, nb = (size_t)1 << log_nb;
    size_t width = (size_t)n_cols + (size_t)n_salt;
    uint64_t *rows = leaves ? leaves : (uint64_t *)malloc(sizeof(uint64_t) * nb * (width ? width : 1));
    if (!rows) return -2;
    int bad = 0;
    /* columns in groups of ORC_COL_GROUP per thread: the transpose then writes ORC_COL_GROUP adjacent words of a
     * row at a time (one cache line) instead of one word per 8 * width bytes */
#define ORC_COL_GROUP 8
    int n_groups = (n_cols + ORC_COL_GROUP - 1) / ORC_COL_GROUP;
#pragma omp parallel
    {
        uint64_t *col = (uint64_t *)malloc(sizeof(uint64_t) * nb * ORC_COL_GROUP);
#pragma omp for schedule(dynamic, 1)
        for (int g = 0; g < n_groups; g++) {
            if (!col) { bad = 1; continue; }
            int c0 = g * ORC_COL_GROUP, k = n_cols - c0 < ORC_COL_GROUP ? n_cols - c0 : ORC_COL_GROUP;
            for (int j = 0; j < k; j++) orc_lde(coeffs + (size_t)(c0 + j) * n, log_n, rate_bits, 7, col + (size_t)j * nb);
            /* transpose + bit-reverse rows: leaf l = evaluation index bitrev(l) */
            for (size_t i = 0; i < nb; i++) {
                uint64_t *row = rows + (size_t)orc_reverse_bits(i, log_nb) * width + c0;
                for (int j = 0; j < k; j++) row[j] = col[(size_t)j * nb + i];
            }
        }
        free(col);
    }
    for (int c = 0; c < n_salt; c++)
        for (size_t l = 0; l < nb; l++)
            rows[l * width + n_cols + c] = salt[(size_t)c * nb + orc_reverse_bits(l, log_nb)];
    int rc = bad ? -2 : orc_merkle_build(rows, log_nb, width, cap_height, digests, cap);
    if (!leaves) free(rows);
    return rc;
}
/* fri/oracle.rs from_values: per-column ifft, then from_coeffs */
int orc_batch_from_values(const uint64_t *values, int log_n, int n_cols, int rate_bits,
                          int cap_height, const uint64_t *salt, int n_salt, uint64_t *coeffs,
                          uint64_t *leaves, uint64_t *digests, uint64_t *cap) {
    if (log_n < 0 || n_cols < 0) return -1;
    size_t n = (size_t)1 << log_n;
    uint64_t *cf = coeffs ? coeffs : (uint64_t *)malloc(sizeof(uint64_t) * n * (n_cols ? n_cols : 1));
    if (!cf) return -2;
    memcpy(cf, values, sizeof(uint64_t) * n * n_cols);
#pragma omp parallel for schedule(dynamic, 1)
    for (int c = 0; c < n_cols; c++) orc_ifft(cf + (size_t)c * n, log_n);
    int rc = orc_batch_from_coeffs(cf, log_n, n_cols, rate_bits, cap_height, salt, n_salt, leaves,
                                   digests, cap);
    if (!coeffs) free(cf);
    return rc;
}

int orc_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
void orc_set_num_threads(int n) {
#ifdef _OPENMP
    omp_set_num_threads(n);
#else
    (void)n;
#endif
}

/* ------------------------------------------------- permutation argument: Zs and partial products
 * plonk/prover.rs wires_permutation_partial_products_and_zs, for one challenge (beta, gamma):
 *   per row i: numerators (w_j + beta k_j x_i + gamma), denominators (w_j + beta sigma_j + gamma), quotients,
 *   quotient_chunk_products(quotients, max_degree)  (util/partial_products.rs: products of chunks of max_degree);
 *   partial_products_and_z_gx(z_x, chunk_products): running products starting from z_x; the last one is Z(g x) and is
 *   swapped with Z(x) so that row i holds [pp_1 .. pp_{num_prods}, Z(x_i)].
 * wires, sigmas: n_routed x n column-major.  out_z[n], out_pp[num_prods][n].  Returns -1 on a zero denominator
 * (upstream panics in batch_multiplicative_inverse). */
int orc_partial_products_and_z(const uint64_t *wires, const uint64_t *sigmas, int log_n, int n_routed, int max_degree,
                               const uint64_t *k_is, uint64_t beta, uint64_t gamma, uint64_t *out_z, uint64_t *out_pp) {
    size_t n = (size_t)1 << log_n;
    int n_chunks = (n_routed + max_degree - 1) / max_degree, num_prods = n_chunks - 1;
    uint64_t g = orc_root_of_unity(log_n), x = 1, z_x = 1;
    for (size_t i = 0; i < n; i++) {
        uint64_t acc = z_x;
        for (int m = 0; m < n_chunks; m++) {
            uint64_t chunk = 1;
            for (int j = m * max_degree; j < n_routed && j < (m + 1) * max_degree; j++) {
                uint64_t w = wires[(size_t)j * n + i] % P;
                uint64_t s_id = mul_(k_is[j] % P, x);
                uint64_t num = add_(add_(w, mul_(beta, s_id)), gamma);
                uint64_t den = add_(add_(w, mul_(beta, sigmas[(size_t)j * n + i] % P)), gamma);
                if (den == 0) return -1;
                chunk = mul_(chunk, mul_(num, orc_inv(den)));
            }
            acc = mul_(acc, chunk);
            if (m < num_prods) out_pp[(size_t)m * n + i] = acc;
        }
        out_z[i] = z_x; /* the swap: Z(x) stays in this row, Z(g x) moves on */
        z_x = acc;
        x = mul_(x, g);
    }
    return 0;
}
