// eval_kernels.cu -- opening-set evaluation: every polynomial of a batch at extension-field points.
// Replaces the `eval_commitment` closure of plonky2 plonk/proof.rs `OpeningSet::new`
// (`c.polynomials.par_iter().map(|p| p.to_extension().eval(z))`, i.e. field/src/polynomial/mod.rs
// `PolynomialCoeffs::eval`, a Horner loop per column) -- SURVEY.md section 8(f) row 2.
//
// Formulation: the exponent of z that multiplies the word at storage index q is linear in the BITS of q
// (weight 2^bit for natural order, 2^(n-1-bit) for the bit-reversed order the commit leaves its
// coefficients in), so  sum_q c_q z^e(q) = sum_q c_q prod_bit m_bit^(q_bit)  with m_bit = z^weight(bit):
// the sum can be folded one index bit at a time, in any bit order, with one multiplier per bit.  That
// lets the loads be coalesced (thread t of a block owns words t + 256 u), the first folds happen in
// registers (base x extension: two multiplies), the next five across lanes by shuffle, three more through
// shared memory, and the rest in a second tiny kernel over per-chunk partial sums.  The coefficients are
// read from HBM exactly once for all points: 8 N C bytes.
#include "internal.h"
#include "gl.cuh"

#define EVAL_THREADS 256
#define EVAL_T_BITS 8
#define EVAL_E_LOG 4  // words per thread in the leaf kernel: chunk = 2^(8 + 4) words

// mt[p][bit] = z_p^weight(bit) for bit < log_n; the rest of the EVAL_MAX_BITS slots stay zero so that folds
// over index bits beyond the polynomial (short inputs) add nothing.
__global__ void eval_mults_kernel(const u64 *points, int n_points, int log_n, int bitrev, u64 *mt) {
    int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n_points) return;
    gl2 z = {gl_canon(points[2 * p]), gl_canon(points[2 * p + 1])};
    u64 *row = mt + (size_t)p * 2 * BSX_EVAL_MAX_BITS;
    for (int j = 0; j < BSX_EVAL_MAX_BITS; j++) row[2 * j] = row[2 * j + 1] = 0;
    for (int j = 0; j < log_n; j++) {
        int bit = bitrev ? log_n - 1 - j : j;
        row[2 * bit] = z.c0;
        row[2 * bit + 1] = z.c1;
        z = gl2_mul(z, z);
    }
}

__device__ __forceinline__ gl2 eval_shfl_xor(gl2 v, int mask) {
    return {__shfl_xor_sync(0xFFFFFFFFu, v.c0, mask), __shfl_xor_sync(0xFFFFFFFFu, v.c1, mask)};
}

// Folds v over the eight thread-index bits; m points at the multiplier of thread bit 0.  Result in thread 0.
__device__ __forceinline__ gl2 eval_block_fold(gl2 v, const u64 *__restrict__ m, gl2 *sm) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < 5; k++) {
        gl2 mk = {m[2 * k], m[2 * k + 1]};
        v = gl2_add(v, gl2_mul(eval_shfl_xor(v, 1 << k), mk));  // meaningful on lanes with bit k clear
    }
    if (lane == 0) sm[warp] = v;
    __syncthreads();
    if (warp == 0) {
        v = sm[lane & 7];
#pragma unroll
        for (int k = 5; k < EVAL_T_BITS; k++) {
            gl2 mk = {m[2 * k], m[2 * k + 1]};
            v = gl2_add(v, gl2_mul(eval_shfl_xor(v, 1 << (k - 5)), mk));
        }
    }
    __syncthreads();
    return v;
}

// grid (n_chunks, n_cols).  partial[(p * n_cols + col) * n_chunks + chunk] = fold of the chunk's 2^12 words.
__global__ void __launch_bounds__(EVAL_THREADS) eval_leaf_kernel(const u64 *__restrict__ coeffs, size_t col_stride,
                                                                 int log_n, const u64 *__restrict__ mt, int n_points,
                                                                 u64 *partial) {
    constexpr int E = 1 << EVAL_E_LOG;
    __shared__ gl2 sm[8];
    const size_t n = (size_t)1 << log_n, base = (size_t)blockIdx.x << (EVAL_T_BITS + EVAL_E_LOG);
    const u64 *__restrict__ src = coeffs + (size_t)blockIdx.y * col_stride;
    u64 c[E];
#pragma unroll
    for (int u = 0; u < E; u++) {
        size_t q = base + threadIdx.x + ((size_t)u << EVAL_T_BITS);
        c[u] = q < n ? __ldcs(src + q) : 0;
    }
    for (int p = 0; p < n_points; p++) {
        const u64 *__restrict__ m = mt + (size_t)p * 2 * BSX_EVAL_MAX_BITS;
        gl2 v[E / 2];
        {
            const u64 m0 = m[2 * EVAL_T_BITS], m1 = m[2 * EVAL_T_BITS + 1];
#pragma unroll
            for (int j = 0; j < E / 2; j++)
                v[j] = {gl_canon(gl_mul_add_lazy(c[2 * j + 1], m0, c[2 * j])), gl_mul(c[2 * j + 1], m1)};
        }
#pragma unroll
        for (int k = 1; k < EVAL_E_LOG; k++) {
            const gl2 mk = {m[2 * (EVAL_T_BITS + k)], m[2 * (EVAL_T_BITS + k) + 1]};
#pragma unroll
            for (int j = 0; j < (E >> (k + 1)); j++) v[j] = gl2_add(v[2 * j], gl2_mul(v[2 * j + 1], mk));
        }
        gl2 r = eval_block_fold(v[0], m, sm);
        if (threadIdx.x == 0) {
            size_t o = ((size_t)p * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
            partial[2 * o] = r.c0;
            partial[2 * o + 1] = r.c1;
        }
    }
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
    const u64 *__restrict__ m = mt + (size_t)p * 2 * BSX_EVAL_MAX_BITS + 2 * bit0;
    gl2 v[4];
#pragma unroll
    for (int u = 0; u < 4; u++) {
        size_t q = threadIdx.x + ((size_t)u << EVAL_T_BITS);
        v[u] = q < g ? gl2{src[2 * q], src[2 * q + 1]} : gl2{0, 0};
    }
    const gl2 m8 = {m[2 * EVAL_T_BITS], m[2 * EVAL_T_BITS + 1]}, m9 = {m[2 * EVAL_T_BITS + 2], m[2 * EVAL_T_BITS + 3]};
    v[0] = gl2_add(v[0], gl2_mul(v[1], m8));
    v[2] = gl2_add(v[2], gl2_mul(v[3], m8));
    v[0] = gl2_add(v[0], gl2_mul(v[2], m9));
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
    return (size_t)n_points * 2 * BSX_EVAL_MAX_BITS + part + part / 2 + 2;
}
// points: n_points extension elements (device).  out: [n_points][n_cols] extension elements (device).
void launch_eval_ext(const u64 *coeffs, size_t col_stride, int log_n, bool bitrev, int n_cols, const u64 *points,
                     int n_points, u64 *work, u64 *out, cudaStream_t st) {
    if (n_cols <= 0 || n_points <= 0) return;
    const size_t chunks = eval_chunks(log_n), cp = (size_t)n_cols * n_points;
    u64 *mt = work, *pa = mt + (size_t)n_points * 2 * BSX_EVAL_MAX_BITS, *pb = pa + 2 * chunks * cp;
    BSX_COUNT_LAUNCH(); eval_mults_kernel<<<1, 32, 0, st>>>(points, n_points, log_n, bitrev ? 1 : 0, mt);
    // the last fold writes straight into out
    u64 *dst = chunks == 1 ? out : pa;
    BSX_COUNT_LAUNCH(); eval_leaf_kernel<<<dim3((unsigned)chunks, (unsigned)n_cols), EVAL_THREADS, 0, st>>>(
        coeffs, col_stride, log_n, mt, n_points, dst);
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
