// ntt_kernels.cu -- batched Goldilocks NTT / iNTT / coset LDE (K2, K3 of SURVEY.md).
// Replaces plonky2 field/src/fft.rs (`fft_dispatch`, `ifft_with_options`) and
// field/src/polynomial/mod.rs (`ifft`, `lde`, `coset_fft_with_options`) as used by
// fri/oracle.rs `PolynomialBatch::from_values` / `lde_values`.
//
// Formulation (not plonky2's bit-reverse + DIT): a size-2^n transform is split n = a + b
// (Cooley-Tukey).  Pass HI runs the 2^a-point decimation-in-frequency sub-transforms over the high
// index bits on tiles of W adjacent columns of the 2^a x 2^b matrix staged in shared memory, then
// multiplies by the inter-pass twiddle omega_N^(i0*k1).  Pass LO runs the 2^b-point sub-transforms
// on contiguous rows.  DIF leaves the output in bit-reversed order, which is exactly the Merkle
// leaf order plonky2 reaches with transpose + reverse_index_bits_in_place, so no permutation pass
// exists.  The LDE's zero padding is never materialised: the 2^r cosets of the big domain are 2^r
// independent size-N transforms of c_i * (g * omega_NB^k)^i, each landing in the contiguous leaf
// block bitrev_r(k).
#include "internal.h"
#include "gl.cuh"

#define NTT_THREADS 512
#define NTT_MAX_A 12
#define NTT_TILE_ELEMS 8192  // 64 KB of shared memory per tile

__device__ __forceinline__ u32 bitrev32(u32 x, int bits) { return bits ? (__brev(x) >> (32 - bits)) : 0; }

struct NttKParams {
    const u64 *src;
    u64 *dst;
    size_t src_stride, dst_stride;
    const u64 *wtab;  // omega_4096^j (or inverse), j < 2048
    const u64 *pre_hi, *pre_lo;
    const u64 *tw_lo, *tw_hi;
    u64 scale;
    int log_n, a, b, w_log;  // tile = 2^a points x 2^w_log adjacent i0
    int src_bitrev, r_log;   // r_log = log2(n_cosets)
    int last;                // HI pass is the only pass: apply scale here
};

// DIF stages over `bits` index bits of a tile in shared memory.
// J_FAST:  sm[p * W + j]  (pass HI: j = adjacent i0, contiguous in global memory)
// !J_FAST: sm[j * R + p]  (pass LO: p contiguous in global memory)
template <bool J_FAST>
__device__ __forceinline__ void dif_stages(u64 *sm, const u64 *__restrict__ wtab, int bits, int w_log) {
    const int R = 1 << bits, W = 1 << w_log;
    const int n_bf = (R >> 1) << w_log;
    for (int s = 0; s < bits; s++) {
        const int h_log = bits - 1 - s;
        for (int bf = threadIdx.x; bf < n_bf; bf += blockDim.x) {
            int j, pp;
            if (J_FAST) {
                j = bf & (W - 1);
                pp = bf >> w_log;
            } else {
                pp = bf & ((R >> 1) - 1);
                j = bf >> (bits - 1);
            }
            int off = pp & ((1 << h_log) - 1), blk = pp >> h_log;
            int p0 = (blk << (h_log + 1)) + off;
            u64 *pu, *pv;
            if (J_FAST) {
                pu = sm + ((size_t)p0 << w_log) + j;
                pv = pu + ((size_t)1 << (h_log + w_log));
            } else {
                pu = sm + ((size_t)j << bits) + p0;
                pv = pu + ((size_t)1 << h_log);
            }
            u64 u = *pu, v = *pv;
            *pu = gl_add(u, v);
            u64 d = gl_sub(u, v);
            // omega_R^(off * 2^s) = omega_4096^(off << (s + 12 - bits))
            *pv = off ? gl_mul(d, wtab[off << (s + NTT_MAX_A - bits)]) : d;
        }
        __syncthreads();
    }
}

// Pass HI: grid (2^b / W, n_cols, n_cosets)
__global__ void __launch_bounds__(NTT_THREADS) ntt_pass_hi(NttKParams P) {
    extern __shared__ u64 sm[];
    const int a = P.a, b = P.b, w_log = P.w_log;
    const int R = 1 << a, W = 1 << w_log;
    const int col = blockIdx.y, k = blockIdx.z;
    const size_t i0_base = (size_t)blockIdx.x << w_log;
    const u64 *src = P.src + (size_t)col * P.src_stride;
    u64 *dst = P.dst + (size_t)col * P.dst_stride + ((size_t)bitrev32(k, P.r_log) << P.log_n);
    const int n_el = R << w_log;

    for (int e = threadIdx.x; e < n_el; e += blockDim.x) {
        int i1, j;
        size_t pos;
        if (P.src_bitrev) {
            // natural index i = i1 * S + i0 lives at bitrev_n(i) = bitrev_b(i0) * R + bitrev_a(i1):
            // W runs of R contiguous words
            j = e >> a;
            int q = e & (R - 1);
            i1 = bitrev32(q, a);
            pos = ((size_t)bitrev32((u32)(i0_base + j), b) << a) + q;
        } else {
            i1 = e >> w_log;
            j = e & (W - 1);
            pos = ((size_t)i1 << b) + i0_base + j;
        }
        u64 v = src[pos];
        if (P.pre_hi) {
            u64 f = gl_mul(P.pre_hi[((size_t)k << a) + i1], P.pre_lo[((size_t)k << b) + i0_base + j]);
            v = gl_mul(v, f);
        }
        sm[((size_t)i1 << w_log) + j] = v;
    }
    __syncthreads();
    dif_stages<true>(sm, P.wtab, a, w_log);
    for (int e = threadIdx.x; e < n_el; e += blockDim.x) {
        int p = e >> w_log, j = e & (W - 1);
        u64 v = sm[e];
        if (b > 0) {
            // position p holds Y[k1], k1 = bitrev_a(p); twiddle omega_N^(i0 * k1)
            u64 ex = (u64)(i0_base + j) * bitrev32(p, a);
            u64 t = P.tw_lo[ex & 4095];
            if (ex >> 12) t = gl_mul(t, P.tw_hi[ex >> 12]);
            v = gl_mul(v, t);
        }
        if (P.last && P.scale != 1) v = gl_mul(v, P.scale);
        dst[((size_t)p << b) + i0_base + j] = v;
    }
}

// Pass LO: in place on rows of 2^b contiguous words; tile = 2^w_log rows.  grid (rows / W, n_cols)
__global__ void __launch_bounds__(NTT_THREADS) ntt_pass_lo(NttKParams P) {
    extern __shared__ u64 sm[];
    const int b = P.b, w_log = P.w_log;
    const int S = 1 << b;
    const int n_el = S << w_log;
    u64 *base = P.dst + (size_t)blockIdx.y * P.dst_stride + ((size_t)blockIdx.x << (b + w_log));
    for (int e = threadIdx.x; e < n_el; e += blockDim.x) sm[e] = base[e];
    __syncthreads();
    dif_stages<false>(sm, P.wtab, b, w_log);
    for (int e = threadIdx.x; e < n_el; e += blockDim.x) {
        u64 v = sm[e];
        if (P.scale != 1) v = gl_mul(v, P.scale);
        base[e] = v;
    }
}

int ntt_max_single_pass_log() { return NTT_MAX_A; }

// per-device opt-in to > 48 KB dynamic shared memory; called from bsx_ctx_create
void ntt_init_device() {
    cudaFuncSetAttribute(ntt_pass_hi, cudaFuncAttributeMaxDynamicSharedMemorySize, 128 * 1024);
    cudaFuncSetAttribute(ntt_pass_lo, cudaFuncAttributeMaxDynamicSharedMemorySize, 128 * 1024);
}

void launch_ntt(const NttTables &T, const NttPlan &plan, const u64 *src, size_t src_stride, u64 *dst,
                size_t dst_stride, int n_cols, cudaStream_t st) {
    if (n_cols <= 0) return;
    NttKParams P;
    P.src = src;
    P.dst = dst;
    P.src_stride = src_stride;
    P.dst_stride = dst_stride;
    P.wtab = plan.inverse ? T.w_inv : T.w_fwd;
    P.pre_hi = plan.pre_hi;
    P.pre_lo = plan.pre_lo;
    P.tw_lo = plan.tw.lo;
    P.tw_hi = plan.tw.hi;
    P.scale = plan.scale;
    P.log_n = plan.log_n;
    P.a = plan.a;
    P.b = plan.b;
    P.src_bitrev = plan.src_bitrev;
    int r_log = 0;
    while ((1 << r_log) < plan.n_cosets) r_log++;
    P.r_log = r_log;
    P.last = plan.b == 0;
    // pass HI: tile of 2^a points x W adjacent i0
    {
        int w_log = 13 - plan.a;  // 8192 elements per tile
        if (w_log < 2) w_log = 2;
        if (w_log > plan.b) w_log = plan.b;
        P.w_log = w_log;
        dim3 grid((unsigned)(((size_t)1 << plan.b) >> w_log), (unsigned)n_cols, (unsigned)plan.n_cosets);
        size_t smem = sizeof(u64) << (plan.a + w_log);
        int threads = NTT_THREADS;
        while (threads > 32 && threads > (1 << (plan.a + w_log))) threads >>= 1;
        BSX_COUNT_LAUNCH(); ntt_pass_hi<<<grid, threads, smem, st>>>(P);
    }
    if (plan.b > 0) {
        size_t rows = ((size_t)plan.n_cosets << plan.log_n) >> plan.b;  // per column
        int w_log = 13 - plan.b;
        if (w_log < 0) w_log = 0;
        while (((size_t)1 << w_log) > rows) w_log--;
        P.w_log = w_log;
        dim3 grid((unsigned)(rows >> w_log), (unsigned)n_cols);
        size_t smem = sizeof(u64) << (plan.b + w_log);
        int threads = NTT_THREADS;
        while (threads > 32 && threads > (1 << (plan.b + w_log))) threads >>= 1;
        BSX_COUNT_LAUNCH(); ntt_pass_lo<<<grid, threads, smem, st>>>(P);
    }
}

__global__ void fill_powers_kernel(u64 *out, u64 base, u64 first, size_t n) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = gl_mul(first, gl_pow(base, i));
}
void launch_fill_powers(u64 *out, u64 base, u64 first, size_t n, cudaStream_t st) {
    if (n == 0) return;
    BSX_COUNT_LAUNCH(); fill_powers_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(out, base, first, n);
}

__global__ void bitrev_copy_kernel(const u64 *__restrict__ src, size_t src_stride, u64 *__restrict__ dst,
                                   size_t dst_stride, int log_n) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >> log_n) return;
    size_t c = blockIdx.y;
    dst[c * dst_stride + i] = src[c * src_stride + bitrev32((u32)i, log_n)];
}
void launch_bitrev_copy(const u64 *src, size_t src_stride, u64 *dst, size_t dst_stride, int log_n, int n_cols,
                        cudaStream_t st) {
    if (n_cols <= 0) return;
    size_t n = (size_t)1 << log_n;
    dim3 grid((unsigned)((n + 255) / 256), (unsigned)n_cols);
    BSX_COUNT_LAUNCH(); bitrev_copy_kernel<<<grid, 256, 0, st>>>(src, src_stride, dst, dst_stride, log_n);
}

// counter-based SplitMix64 stream reduced mod p: salt columns when the caller supplies none
__global__ void fill_random_kernel(u64 *out, size_t n, u64 seed) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    u64 z = seed + (i + 1) * 0x9E3779B97F4A7C15ULL;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    z ^= z >> 31;
    out[i] = z % GL_P;
}
void launch_fill_random(u64 *out, size_t n, u64 seed, cudaStream_t st) {
    if (n == 0) return;
    BSX_COUNT_LAUNCH(); fill_random_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(out, n, seed);
}
