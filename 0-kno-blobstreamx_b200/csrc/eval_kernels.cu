// eval_kernels.cu -- opening-set evaluation: every polynomial of a batch at extension-field points.
// Replaces the `eval_commitment` closure of plonky2 plonk/proof.rs `OpeningSet::new`
// (`c.polynomials.par_iter().map(|p| p.to_extension().eval(z))`, i.e. field/src/polynomial/mod.rs
// `PolynomialCoeffs::eval`, a Horner loop per column) -- SURVEY.md section 8(f) row 2.
//
// Formulation: the exponent of z that multiplies the word at storage index q is linear in the BITS of q
// (weight 2^bit for natural order, 2^(n-1-bit) for the bit-reversed order the commit leaves its
// coefficients in), so  sum_q c_q z^e(q) = sum_q c_q prod_bit m_bit^(q_bit)  with m_bit = z^weight(bit):
// the sum can be folded one index bit at a time, in any bit order, with one multiplier per bit.  That
// lets the loads be coalesced (thread t of a block owns words t + 256 u).  The fold over u is a dot product of
// base-field words with 64 precomputed extension weights (two 8-instruction column-accumulator MACs per
// word, no reduction until the end); the eight thread bits follow by shuffle and shared memory, and the
// per-chunk partial sums are folded by a second tiny kernel.  Two points share one pass, so the opening
// set reads its coefficients from HBM once: 8 N C bytes.
#include "internal.h"
#include "gl.cuh"

#define EVAL_THREADS 256
#define EVAL_T_BITS 8
#define EVAL_E_LOG 6  // words per thread in the leaf kernel: chunk = 2^(8 + 6) words
#define EVAL_E (1 << EVAL_E_LOG)
// per point: BSX_EVAL_MAX_BITS fold multipliers {m.c0, m.c1, 7 * m.c1}, then EVAL_E in-thread weights {w.c0, w.c1}
#define EVAL_ROW_WORDS (3 * BSX_EVAL_MAX_BITS + 2 * EVAL_E)

// m[bit] = z^weight(bit) for bit < log_n; the remaining slots stay zero so that folds over index bits beyond
// the polynomial (short inputs) add nothing.  w[u] = prod over the set bits k of u of m[8 + k].
__global__ void eval_mults_kernel(const u64 *points, int n_points, int log_n, int bitrev, u64 *mt) {
    int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n_points) return;
    gl2 z = {gl_canon(points[2 * p]), gl_canon(points[2 * p + 1])};
    u64 *row = mt + (size_t)p * EVAL_ROW_WORDS, *w = row + 3 * BSX_EVAL_MAX_BITS;
    for (int j = 0; j < 3 * BSX_EVAL_MAX_BITS; j++) row[j] = 0;
    for (int j = 0; j < log_n; j++) {
        int bit = bitrev ? log_n - 1 - j : j;
        row[3 * bit] = z.c0;
        row[3 * bit + 1] = z.c1;
        row[3 * bit + 2] = gl_mul_small(z.c1, 7);
        z = gl2_mul(z, z);
    }
    w[0] = 1;
    w[1] = 0;
    for (int k = 0; k < EVAL_E_LOG; k++) {
        gl2 mk = {row[3 * (EVAL_T_BITS + k)], row[3 * (EVAL_T_BITS + k) + 1]};
        for (int u = 0; u < (1 << k); u++) {
            gl2 r = gl2_mul(gl2{w[2 * u], w[2 * u + 1]}, mk);
            w[2 * (u + (1 << k))] = r.c0;
            w[2 * (u + (1 << k)) + 1] = r.c1;
        }
    }
}

__device__ __forceinline__ gl2 eval_shfl_xor(gl2 v, int mask) {
    return {__shfl_xor_sync(0xFFFFFFFFu, v.c0, mask), __shfl_xor_sync(0xFFFFFFFFu, v.c1, mask)};
}
// v + o * m with m given as {m0, m1, 7 m1}; v, o canonical -> canonical
__device__ __forceinline__ gl2 eval_fma(gl2 v, gl2 o, const u64 *__restrict__ m) {
    const u64 m0 = m[0], m1 = m[1], m7 = m[2];
    u64 t0 = gl_mul_add_lazy(o.c1, m7, gl_mul_lazy(o.c0, m0));
    u64 t1 = gl_mul_add_lazy(o.c1, m0, gl_mul_lazy(o.c0, m1));
    return {gl_canon_fast(gl_add_lazy_canon(t0, v.c0)), gl_canon_fast(gl_add_lazy_canon(t1, v.c1))};
}

// Folds v over the eight thread-index bits; m points at the multiplier of thread bit 0.  Result in thread 0.
__device__ __forceinline__ gl2 eval_block_fold(gl2 v, const u64 *__restrict__ m, gl2 *sm) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < 5; k++) v = eval_fma(v, eval_shfl_xor(v, 1 << k), m + 3 * k);  // meaningful where bit k is clear
    if (lane == 0) sm[warp] = v;
    __syncthreads();
    if (warp == 0) {
        v = sm[lane & 7];
#pragma unroll
        for (int k = 5; k < EVAL_T_BITS; k++) v = eval_fma(v, eval_shfl_xor(v, 1 << (k - 5)), m + 3 * k);
    }
    __syncthreads();
    return v;
}

// grid (n_chunks, n_cols).  partial[(p * n_cols + col) * n_chunks + chunk] = fold of the chunk's 2^14 words.
// A thread owns words t + 256 u: the fold over u is a dot product of base-field words with the extension weights
// w[u], two column-accumulator MACs (8 instructions each) per word and point.
template <int NP>
__global__ void __launch_bounds__(EVAL_THREADS) eval_leaf_kernel(const u64 *__restrict__ coeffs, size_t col_stride,
                                                                 int log_n, const u64 *__restrict__ mt, int p_total,
                                                                 int p_begin, u64 *partial) {
    __shared__ gl2 sm[8];
    __shared__ u64 ws[NP][2 * EVAL_E];
    const size_t n = (size_t)1 << log_n, base = (size_t)blockIdx.x << (EVAL_T_BITS + EVAL_E_LOG);
    const u64 *__restrict__ src = coeffs + (size_t)blockIdx.y * col_stride + base + threadIdx.x;
    for (int i = threadIdx.x; i < NP * 2 * EVAL_E; i += blockDim.x)
        ws[i / (2 * EVAL_E)][i % (2 * EVAL_E)] =
            mt[(size_t)(p_begin + i / (2 * EVAL_E)) * EVAL_ROW_WORDS + 3 * BSX_EVAL_MAX_BITS + i % (2 * EVAL_E)];
    __syncthreads();
    const size_t left = n - base;  // words of this column from the chunk start
    const int u_full = left >= ((size_t)EVAL_E << EVAL_T_BITS) ? EVAL_E : (int)(left >> EVAL_T_BITS);
    gl_acc3 A[NP][2];
#pragma unroll
    for (int p = 0; p < NP; p++) A[p][0] = A[p][1] = gl_acc3_zero();
#pragma unroll 8
    for (int u = 0; u < u_full; u++) {
        const u64 c = __ldcs(src + ((size_t)u << EVAL_T_BITS));
#pragma unroll
        for (int p = 0; p < NP; p++) {
            gl_acc3_mac(A[p][0], c, ws[p][2 * u]);
            gl_acc3_mac(A[p][1], c, ws[p][2 * u + 1]);
        }
    }
    if (u_full < EVAL_E && threadIdx.x < (left & (EVAL_THREADS - 1))) {  // polynomials shorter than 256 words
        const u64 c = src[(size_t)u_full << EVAL_T_BITS];
#pragma unroll
        for (int p = 0; p < NP; p++) {
            gl_acc3_mac(A[p][0], c, ws[p][2 * u_full]);
            gl_acc3_mac(A[p][1], c, ws[p][2 * u_full + 1]);
        }
    }
#pragma unroll
    for (int p = 0; p < NP; p++) {
        gl2 v = {gl_canon_fast(gl_acc3_reduce(A[p][0])), gl_canon_fast(gl_acc3_reduce(A[p][1]))};
        gl2 r = eval_block_fold(v, mt + (size_t)(p_begin + p) * EVAL_ROW_WORDS, sm);
        if (threadIdx.x == 0) {
            size_t o = ((size_t)(p_begin + p) * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
            partial[2 * o] = r.c0;
            partial[2 * o + 1] = r.c1;
        }
    }
    (void)p_total;
}

// grid (m_in >> g_log, n_points * n_cols): folds groups of 2^g_log <= 1024 partial sums over index bits
// bit0 .. bit0 + g_log - 1 of the original storage index.
__global__ void __launch_bounds__(EVAL_THREADS) eval_fold_kernel(const u64 *__restrict__ in, size_t m_in, int g_log,
                                                                 int bit0, const u64 *__restrict__ mt, int n_cols,
                                                                 u64 *out) {
    __shared__ gl2 sm[8];
    const int cp = blockIdx.y, p = cp / n_cols;
    const size_t g = (size_t)1 << g_log;
    const u64 *__restrict__ src = in + 2 * ((size_t)cp * m_in + (size_t)blockIdx.x * g);
    const u64 *__restrict__ m = mt + (size_t)p * EVAL_ROW_WORDS + 3 * bit0;
    gl2 v[4];
#pragma unroll
    for (int u = 0; u < 4; u++) {
        size_t q = threadIdx.x + ((size_t)u << EVAL_T_BITS);
        v[u] = q < g ? gl2{src[2 * q], src[2 * q + 1]} : gl2{0, 0};
    }
    v[0] = eval_fma(v[0], v[1], m + 3 * EVAL_T_BITS);
    v[2] = eval_fma(v[2], v[3], m + 3 * EVAL_T_BITS);
    v[0] = eval_fma(v[0], v[2], m + 3 * (EVAL_T_BITS + 1));
    gl2 r = eval_block_fold(v[0], m, sm);
    if (threadIdx.x == 0) {
        size_t o = (size_t)cp * gridDim.x + blockIdx.x;
        out[2 * o] = r.c0;
        out[2 * o + 1] = r.c1;
    }
}

size_t eval_chunks(int log_n) {
    const int cl = EVAL_T_BITS + EVAL_E_LOG;
    return log_n > cl ? (size_t)1 << (log_n - cl) : 1;
}
// words of workspace: multiplier table + two partial-sum buffers
size_t eval_workspace_words(int log_n, int n_cols, int n_points) {
    size_t part = 2 * eval_chunks(log_n) * (size_t)n_cols * (size_t)n_points;
    return (size_t)n_points * EVAL_ROW_WORDS + part + part / 2 + 2;
}
// points: n_points extension elements (device).  out: [n_points][n_cols] extension elements (device).
void launch_eval_ext(const u64 *coeffs, size_t col_stride, int log_n, bool bitrev, int n_cols, const u64 *points,
                     int n_points, u64 *work, u64 *out, cudaStream_t st) {
    if (n_cols <= 0 || n_points <= 0) return;
    const size_t chunks = eval_chunks(log_n), cp = (size_t)n_cols * n_points;
    u64 *mt = work, *pa = mt + (size_t)n_points * EVAL_ROW_WORDS, *pb = pa + 2 * chunks * cp;
    BSX_COUNT_LAUNCH(); eval_mults_kernel<<<1, 32, 0, st>>>(points, n_points, log_n, bitrev ? 1 : 0, mt);
    // the last fold writes straight into out
    u64 *dst = chunks == 1 ? out : pa;
    const dim3 grid((unsigned)chunks, (unsigned)n_cols);
    for (int p = 0; p < n_points; p += 2) {  // two points share one pass over the coefficients
        BSX_COUNT_LAUNCH();
        if (p + 1 < n_points)
            eval_leaf_kernel<2><<<grid, EVAL_THREADS, 0, st>>>(coeffs, col_stride, log_n, mt, n_points, p, dst);
        else
            eval_leaf_kernel<1><<<grid, EVAL_THREADS, 0, st>>>(coeffs, col_stride, log_n, mt, n_points, p, dst);
    }
    size_t m = chunks;
    int bit0 = EVAL_T_BITS + EVAL_E_LOG;
    u64 *cur = pa, *nxt = pb;
    while (m > 1) {
        int ml = 0;
        while (((size_t)1 << ml) < m) ml++;
        int g_log = ml < 10 ? ml : 10;
        size_t m_out = m >> g_log;
        u64 *o = m_out == 1 ? out : nxt;
        BSX_COUNT_LAUNCH(); eval_fold_kernel<<<dim3((unsigned)m_out, (unsigned)cp), EVAL_THREADS, 0, st>>>(
            cur, m, g_log, bit0, mt, n_cols, o);
        bit0 += g_log;
        m = m_out;
        u64 *t = cur; cur = nxt; nxt = t;
    }
}
