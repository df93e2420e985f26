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
// block bitrev_r(k); one CTA produces all 2^r cosets of its tile, so the coefficients are read from
// HBM once (the other 2^r - 1 reads hit L2).
//
// Inside a tile the stages run in groups of up to four: a thread pulls 2^rho points out of shared
// memory, runs rho butterfly stages on them in registers (12-instruction canonical add/sub pair,
// 20-instruction multiply) and puts them back, so a 2^10-point sub-transform makes three trips
// through shared memory instead of ten.  Pass HI's tile is padded by one word in 32 so that the
// stride-2^rho accesses of the last group and the bit-reversed coefficient scatter stay conflict-free;
// pass LO's tile is one contiguous run of global memory, so it comes in and goes out as TMA bulk tensor
// copies with the 128-byte hardware swizzle doing what the pad words do (ntt_pass_lo_tma).  The kernels
// are bound by integer issue (ALU carry chains + IMAD.WIDE), not by HBM: DESIGN.md section 3.4, profiles/.
#include "internal.h"
// pipe-balanced multiply + canonicalisation for the butterflies (gl.cuh gl_mul_canon): the transforms are ALU-pipe bound
#ifndef GL_FMA_MUL
#define GL_FMA_MUL 1
#endif
#ifndef GL_FMA_CANON
#define GL_FMA_CANON 1
#endif
#include "gl.cuh"

// 1: pass LO moves its tile with bulk tensor copies (TMA) instead of LDG/STS + LDS/STG loops: LDE 20.69 -> 19.77 ms,
// iNTT 2.69 -> 2.56 ms at 2^20 x 135 (profiles/r02_ntt_variants.txt); 0 keeps the loops (A/B builds)
#ifndef NTT_LO_TMA
#define NTT_LO_TMA 1
#endif
#if NTT_LO_TMA
#include <cuda.h>
#endif

#define NTT_THREADS 256
#define NTT_MAX_A 12
#ifndef NTT_W_MIN_A12
#define NTT_W_MIN_A12 1
#endif

__device__ __forceinline__ u32 bitrev32(u32 x, int bits) { return bits ? (__brev(x) >> (32 - bits)) : 0; }

struct NttGroups {
    int n;
    int rho[4];
};

struct NttKParams {
    const u64 *src;
    u64 *dst;
    size_t src_stride, dst_stride;
    const u64 *wtab;  // omega_4096^j (or inverse), j < 2048
    const u64 *pre_hi, *pre_lo;
    const u64 *tw_lo, *tw_hi;
    const u64 *post_full;  // [n_cosets][2^a][2^b]: scale * pre_lo * inter-pass twiddle, or null
    u64 scale;
    int log_n, a, b, w_log;  // tile = 2^a points x 2^w_log adjacent i0
    int src_bitrev, r_log;   // r_log = log2(n_cosets)
    int pitch;               // words between tile columns in shared memory
    NttGroups groups;
    int n_dest, dest_lc;     // row-sharded output (see NttPlan)
    int dest_tma;            // the destination slabs have tensor maps (ntt_pass_lo_tma)
    u64 *dest[BSX_MAX_RANKS];
};

// final position of leaf l of column col: local column-major buffer, or the owning rank's row slab
__device__ __forceinline__ u64 *ntt_out(const NttKParams &P, int col, size_t l) {
    if (P.n_dest == 0) return P.dst + (size_t)col * P.dst_stride + l;
    return P.dest[l >> P.dest_lc] + ((size_t)col << P.dest_lc) + (l & (((size_t)1 << P.dest_lc) - 1));
}

__device__ __forceinline__ u64 gl_mul_c(u64 a, u64 b) { return gl_mul_canon(a, b); }

// Tile layout in shared memory: W columns of 2^bits points, column j at j * pitch, point p at
// p + (p >> 5) + (p >> 9).  The padding keeps the stride-2^rho accesses of the last stage group and the
// bit-reversed scatter of the coefficient load conflict-free; pitch == 2 (mod 16) does the same for
// the column-interleaved global loads / stores of pass HI.
__device__ __forceinline__ int ntt_swz(int p) { return p + (p >> 5) + (p >> 9); }
__device__ __forceinline__ int ntt_sm(int j, int p, int pitch) { return j * pitch + ntt_swz(p); }
// Twiddle table position: stage s reads entries e = off << s, a stride-2^s walk across the lanes; XOR-ing
// bits 5.. into the low bits spreads every such walk over all banks.
__device__ __forceinline__ int ntt_tw_pos(int e) { return e ^ (e >> 5); }
// Unpadded tile as a bulk tensor copy with CU_TENSOR_MAP_SWIZZLE_128B leaves it (NTT_LO_TMA): the 16-byte chunk index
// inside a 128-byte line is XOR-ed with the line index mod 8, i.e. bits 1..3 of the word index with bits 4..6.
__device__ __forceinline__ int ntt_xsw(int p) { return p ^ (((p >> 4) & 7) << 1); }
template <bool XSW>
__device__ __forceinline__ int ntt_pos(int p) { return XSW ? ntt_xsw(p) : ntt_swz(p); }
#define NTT_IO_UNROLL 8
#ifndef NTT_LO_VEC
#define NTT_LO_VEC 1
#endif
#ifndef NTT_HI_VEC
#define NTT_HI_VEC 1
#endif
// 1: the last stage group multiplies by its twiddles (powers of omega_16 = -2^60) with shifts (gl_mul_pow2_canon)
#ifndef NTT_SHIFT_TW
#define NTT_SHIFT_TW 1
#endif
// 1: the other stage groups run as radix-2^rho steps: shift twiddles inside, and the factors that depend on the thread's
// position, omega^((lo * bitrev(m)) << s0), collected into ONE multiplication per output after the group (rho * 2^rho / 2
// table multiplications become 2^rho - 1).  Needs the full table of 2^bits twiddles in shared memory.
#ifndef NTT_RADIX_GROUPS
#define NTT_RADIX_GROUPS 1
#endif
#define NTT_TW_FULL (NTT_SHIFT_TW && NTT_RADIX_GROUPS)
// Twiddle table in shared memory: all 2^bits powers and radix groups (FULLTW kernels) for tiles of up to 2^10 points; 2^11-
// and 2^12-point tiles keep the half table and the per-stage twiddles -- a full table would cost them the third CTA per
// SM (measured: 2^24-row iNTT 7.28 -> 7.95 ms), and folding omega^(N/2) = -1 into the operand instead was slower too (7.81).
static inline bool ntt_tw_full(int bits) { return NTT_TW_FULL && bits < 11; }
#ifndef NTT_HI_UNROLL
#define NTT_HI_UNROLL 4
#endif
#define NTT_PRAGMA(x) _Pragma(#x)
#define NTT_UNROLL(n) NTT_PRAGMA(unroll n)

// One group of RHO DIF stages, starting at stage s0 of a 2^bits-point transform.  A work item owns
// the 2^RHO points p = ((hi << RHO | m) << lobits) | lo of one column.  Butterfly (m0, m0 + half) of
// local stage t has twiddle omega_{2^(bits - s0 - t)}^(mlow << lobits | lo)
//   = omega_{2^bits}^((mlow << (bits - RHO + t)) + (lo << (s0 + t))), read from the table in shared memory.
template <int RHO, bool LAST, bool INV, bool FULLTW, bool XSW = false>
__device__ __forceinline__ void dif_group(u64 *sm, const u64 *__restrict__ tws, int bits, int w_log, int s0,
                                          int pitch) {
    constexpr int R = 1 << RHO;
    const int lobits = LAST ? 0 : bits - s0 - RHO;
    const int n_items = 1 << (bits + w_log - RHO);
    for (int it = threadIdx.x; it < n_items; it += blockDim.x) {
        const int lo = it & ((1 << lobits) - 1);
        const int r = it >> lobits;
        const int hi = r & ((1 << s0) - 1);
        const int j = r >> s0;
        const int p0 = (hi << (RHO + lobits)) | lo;
        u64 *col = sm + j * pitch;
        u64 x[R];
#pragma unroll
        for (int m = 0; m < R; m++) {
            int p = p0 + (m << lobits);
            x[m] = col[ntt_pos<XSW>(p)];
        }
#pragma unroll
        for (int t = 0; t < RHO; t++) {
            const int half = R >> (t + 1);
            const int e_lo = LAST ? 0 : (lo << (s0 + t));
#pragma unroll
            for (int bf = 0; bf < R / 2; bf++) {
                const int mlow = bf & (half - 1);
                const int m0 = ((bf / half) * 2 * half) + mlow;
                const int m1 = m0 + half;
#if NTT_SHIFT_TW
                if (LAST || FULLTW) {
                    // twiddle omega_{2^RHO}^(mlow << t) = omega_16^j = 2^(156 j) (inverse: 2^(-156 j)); 2^96 == -1, and a
                    // negative twiddle is the subtraction taken the other way round: (a - b) (-w) = (b - a) w
                    const int j = (mlow << t) << (4 - RHO);
                    const int e = INV ? (192 - (156 * j) % 192) % 192 : (156 * j) % 192;
                    const bool neg = e >= 96;
                    const u64 a0 = x[m0], b0 = x[m1];
                    x[m0] = gl_sub_c(a0, gl_neg_p(b0));
                    const u64 d = neg ? gl_sub_c(b0, a0) : gl_sub_c(a0, b0);
                    x[m1] = mlow != 0 ? gl_mul_pow2_canon(d, neg ? e - 96 : e) : d;
                    continue;
                }
#endif
                gl_butterfly_c(x[m0], x[m1]);
                if (!LAST || mlow != 0) x[m1] = gl_mul_c(x[m1], tws[ntt_tw_pos(e_lo | (mlow << (bits - RHO + t)))]);
            }
        }
        if (!LAST && FULLTW) {
            // position m holds output k = bitrev_RHO(m) of the radix-R step: its factor is omega_{2^bits}^((lo k) << s0)
            const int base = lo << s0;
#pragma unroll
            for (int m = 1; m < R; m++) {
                const int k = (int)(__brev((unsigned)m) >> (32 - RHO));
                x[m] = gl_mul_c(x[m], tws[ntt_tw_pos(base * k)]);
            }
        }
#pragma unroll
        for (int m = 0; m < R; m++) {
            int p = p0 + (m << lobits);
            col[ntt_pos<XSW>(p)] = x[m];
        }
    }
}

template <bool INV, bool FULLTW, bool XSW = false>
__device__ __forceinline__ void dif_tile(u64 *sm, const u64 *__restrict__ tws, int bits, int w_log, int pitch,
                                         const NttGroups &G) {
    int s0 = 0;
    for (int g = 0; g < G.n; g++) {
        const int rho = G.rho[g];
        if (g == G.n - 1) {
            switch (rho) {
                case 1: dif_group<1, true, INV, FULLTW, XSW>(sm, tws, bits, w_log, s0, pitch); break;
                case 2: dif_group<2, true, INV, FULLTW, XSW>(sm, tws, bits, w_log, s0, pitch); break;
                case 3: dif_group<3, true, INV, FULLTW, XSW>(sm, tws, bits, w_log, s0, pitch); break;
                default: dif_group<4, true, INV, FULLTW, XSW>(sm, tws, bits, w_log, s0, pitch); break;
            }
        } else {
            switch (rho) {
                case 1: dif_group<1, false, INV, FULLTW, XSW>(sm, tws, bits, w_log, s0, pitch); break;
                case 2: dif_group<2, false, INV, FULLTW, XSW>(sm, tws, bits, w_log, s0, pitch); break;
                case 3: dif_group<3, false, INV, FULLTW, XSW>(sm, tws, bits, w_log, s0, pitch); break;
                default: dif_group<4, false, INV, FULLTW, XSW>(sm, tws, bits, w_log, s0, pitch); break;
            }
        }
        __syncthreads();
        s0 += rho;
    }
}

// omega_{2^bits}^e, e < 2^(bits-1) (NTT_TW_FULL: e < 2^bits), from the omega_4096 table into shared memory
__device__ __forceinline__ void load_twiddles(u64 *tws, const u64 *__restrict__ wtab, int bits, bool full) {
    const int n = full ? (1 << bits) : (bits > 0 ? 1 << (bits - 1) : 0);
    for (int e = threadIdx.x; e < n; e += blockDim.x)
        tws[ntt_tw_pos(e)] = __ldg(wtab + ((size_t)e << (NTT_MAX_A - bits)));
}

#ifdef NTT_PROFILE
// debug build only (make EXTRA=-DNTT_PROFILE): per-phase SM cycles of pass HI summed over CTAs and cosets
__device__ unsigned long long ntt_prof[4];
extern "C" void bsx_debug_ntt_profile(unsigned long long *out, int reset) {
    cudaMemcpyFromSymbol(out, ntt_prof, sizeof(ntt_prof));
    if (reset) {
        unsigned long long z[4] = {0, 0, 0, 0};
        cudaMemcpyToSymbol(ntt_prof, z, sizeof(z));
    }
}
#define NTT_PROF_MARK(i)                                              \
    do {                                                              \
        __syncthreads();                                              \
        if (threadIdx.x == 0) {                                       \
            long long t_now = clock64();                              \
            atomicAdd(&ntt_prof[i], (unsigned long long)(t_now - t_prev)); \
            t_prev = t_now;                                           \
        }                                                             \
    } while (0)
#else
#define NTT_PROF_MARK(i)
#endif

// Pass HI: grid (n_cols, 2^b / W); the CTA loops over the 2^r_log cosets of its tile.
// The source order, the presence of coset factors and the kind of output factor are template parameters so
// that the load / store loops are branch-free and their global loads can be issued NTT_HI_UNROLL deep
// (with the conditions as run-time branches ptxas kept one load in flight per thread and the pass spent
// 72 % of its time in these two loops).  POST: 0 = none (single pass; optional scale), 1 = two-level
// inter-pass twiddle, 2 = full post-factor table.
template <bool SRC_BITREV, bool HAS_PRE, int POST, bool INV, bool FULLTW>
__global__ void __launch_bounds__(NTT_THREADS, 3) ntt_pass_hi(NttKParams P) {
    extern __shared__ u64 sm_all[];
    const int a = P.a, b = P.b, w_log = P.w_log, pitch = P.pitch;
    const int R = 1 << a, W = 1 << w_log;
    u64 *tws = sm_all, *sm = sm_all + (FULLTW ? R : (R >> 1));
    const int col = blockIdx.x;
    const size_t i0_base = (size_t)blockIdx.y << w_log;
    const u64 *__restrict__ src = P.src + (size_t)col * P.src_stride;
    const int n_el = R << w_log;
    const int n_cosets = 1 << P.r_log;
    load_twiddles(tws, P.wtab, a, FULLTW);
#ifdef NTT_PROFILE
    long long t_prev = clock64();
#endif

    for (int k = 0; k < n_cosets; k++) {
        const size_t blk = (size_t)bitrev32(k, P.r_log) << P.log_n;
        u64 *dst = P.dst + (size_t)col * P.dst_stride + blk;
        const bool remap = P.n_dest != 0 && POST == 0;
        // bit-reversed source: index the bit-reversed copy of the factor row by q (contiguous across lanes)
        const u64 *__restrict__ pre_hi = HAS_PRE ? P.pre_hi + ((size_t)(SRC_BITREV ? n_cosets + k : k) << a) : nullptr;
        const u64 *__restrict__ pre_lo = HAS_PRE ? P.pre_lo + ((size_t)k << b) + i0_base : nullptr;
        // LDE from bit-reversed coefficients (the from_values path): column j of the tile is one contiguous run
        // of R words, and the coset factor depends only on the position q in the run.  A thread owns q = tid +
        // t * blockDim (t < R / blockDim <= 4): factors and shared-memory offsets are computed once per coset and
        // reused for all W columns, and each column costs one coalesced load, one multiply and one store per word.
        const bool run_path = SRC_BITREV && HAS_PRE && POST == 2 && R >= (int)blockDim.x && R <= 4 * (int)blockDim.x;
        if (run_path) {
            const int qt = R / (int)blockDim.x;
            u64 fq[4];
            int sw[4];
#pragma unroll
            for (int t = 0; t < 4; t++) {
                const int q = threadIdx.x + (t < qt ? t : 0) * blockDim.x;
                fq[t] = __ldg(pre_hi + q);
                sw[t] = ntt_swz(bitrev32(q, a));
            }
            for (int j = 0; j < W; j += 2) {
                const u64 *__restrict__ row0 = src + ((size_t)bitrev32((u32)(i0_base + j), b) << a) + threadIdx.x;
                const u64 *__restrict__ row1 = src + ((size_t)bitrev32((u32)(i0_base + j + 1), b) << a) + threadIdx.x;
                u64 v0[4], v1[4];
#pragma unroll
                for (int t = 0; t < 4; t++)
                    if (t < qt) {
                        v0[t] = row0[t * blockDim.x];
                        v1[t] = row1[t * blockDim.x];
                    }
#pragma unroll
                for (int t = 0; t < 4; t++)
                    if (t < qt) {
                        sm[j * pitch + sw[t]] = gl_mul_c(v0[t], fq[t]);
                        sm[(j + 1) * pitch + sw[t]] = gl_mul_c(v1[t], fq[t]);
                    }
            }
        } else
        for (int e0 = threadIdx.x; e0 < n_el; e0 += blockDim.x * NTT_HI_UNROLL) {
            u64 v[NTT_HI_UNROLL], f[NTT_HI_UNROLL];
            int si[NTT_HI_UNROLL];
#pragma unroll
            for (int u = 0; u < NTT_HI_UNROLL; u++) {
                const int e_raw = e0 + u * blockDim.x;
                const bool ok = e_raw < n_el;
                const int e = ok ? e_raw : e0;  // out-of-range slots re-read a valid word and are not stored
                int i1, j;
                size_t pos;
                if (SRC_BITREV) {
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
                v[u] = src[pos];
                if (HAS_PRE) {
                    f[u] = __ldg(pre_hi + (SRC_BITREV ? (e & (R - 1)) : i1));
                    if (POST != 2) f[u] = gl_mul_lazy(f[u], __ldg(pre_lo + j));
                }
                si[u] = ok ? ntt_sm(j, i1, pitch) : -1;
            }
#pragma unroll
            for (int u = 0; u < NTT_HI_UNROLL; u++)
                if (si[u] >= 0) sm[si[u]] = HAS_PRE ? gl_mul_c(v[u], f[u]) : gl_canon_fast(v[u]);
        }
        __syncthreads();
        NTT_PROF_MARK(0);
        dif_tile<INV, FULLTW>(sm, tws, a, w_log, pitch, P.groups);
        NTT_PROF_MARK(1);
        const u64 *__restrict__ post = POST == 2 ? P.post_full + ((size_t)k << P.log_n) + i0_base : nullptr;
#if NTT_HI_VEC
        // 128-bit post-factor loads and output stores: a thread takes the pair of adjacent tile columns (j, j + 1) of one
        // row p, i.e. 16 contiguous bytes of the factor table and of the output
        if (POST == 2 && w_log >= 1 && !remap) {
            const int half = n_el >> 1;
            for (int q0 = threadIdx.x; q0 < half; q0 += blockDim.x * NTT_HI_UNROLL) {
                u64 v0[NTT_HI_UNROLL], v1[NTT_HI_UNROLL];
                ulonglong2 f[NTT_HI_UNROLL];
                size_t oo[NTT_HI_UNROLL];
#pragma unroll
                for (int u = 0; u < NTT_HI_UNROLL; u++) {
                    const int q_raw = q0 + u * blockDim.x;
                    const int q = q_raw < half ? q_raw : q0;
                    const int p = q >> (w_log - 1), j = (q & ((W >> 1) - 1)) << 1;
                    v0[u] = sm[ntt_sm(j, p, pitch)];
                    v1[u] = sm[ntt_sm(j + 1, p, pitch)];
                    oo[u] = ((size_t)p << b) + j;
                    f[u] = __ldg(reinterpret_cast<const ulonglong2 *>(post + oo[u]));
                }
#pragma unroll
                for (int u = 0; u < NTT_HI_UNROLL; u++) {
                    if (q0 + u * blockDim.x >= half) continue;
                    __stcs(reinterpret_cast<ulonglong2 *>(dst + i0_base + oo[u]),
                           make_ulonglong2(gl_mul_c(v0[u], f[u].x), gl_mul_c(v1[u], f[u].y)));
                }
            }
        } else
#endif
        for (int e0 = threadIdx.x; e0 < n_el; e0 += blockDim.x * NTT_HI_UNROLL) {
            u64 v[NTT_HI_UNROLL], f[NTT_HI_UNROLL];
            size_t oo[NTT_HI_UNROLL];
#pragma unroll
            for (int u = 0; u < NTT_HI_UNROLL; u++) {
                const int e_raw = e0 + u * blockDim.x;
                const int e = e_raw < n_el ? e_raw : e0;
                const int p = e >> w_log, j = e & (W - 1);
                v[u] = sm[ntt_sm(j, p, pitch)];
                oo[u] = ((size_t)p << b) + j;
                if (POST == 2) {
                    f[u] = __ldg(post + oo[u]);
                } else if (POST == 1) {
                    // position p holds Y[k1], k1 = bitrev_a(p); twiddle omega_N^(i0 * k1)
                    u64 ex = (u64)(i0_base + j) * bitrev32(p, a);
                    f[u] = __ldg(P.tw_lo + (ex & 4095));
                    if (ex >> 12) f[u] = gl_mul_lazy(f[u], __ldg(P.tw_hi + (ex >> 12)));
                } else {
                    f[u] = P.scale;
                }
            }
#pragma unroll
            for (int u = 0; u < NTT_HI_UNROLL; u++) {
                if (e0 + u * blockDim.x >= n_el) continue;
                u64 r = (POST != 0 || P.scale != 1) ? gl_mul_c(v[u], f[u]) : v[u];
                // streaming stores: the LDE output is not read again before the whole pass is over
                if (POST == 0 && remap) __stcs(ntt_out(P, col, blk + i0_base + oo[u]), r);
                else __stcs(dst + i0_base + oo[u], r);
            }
        }
        __syncthreads();
        NTT_PROF_MARK(2);
    }
}

typedef void (*ntt_hi_fn)(NttKParams);
#define NTT_HI_ROW(B, H, I, F) {ntt_pass_hi<B, H, 0, I, F>, ntt_pass_hi<B, H, 1, I, F>, ntt_pass_hi<B, H, 2, I, F>}
#define NTT_HI_DIR(I, F) {{NTT_HI_ROW(false, false, I, F), NTT_HI_ROW(false, true, I, F)}, {NTT_HI_ROW(true, false, I, F), NTT_HI_ROW(true, true, I, F)}}
static ntt_hi_fn ntt_hi_variant(bool src_bitrev, bool has_pre, int post, bool inverse, bool full_tw) {
    static const ntt_hi_fn tab[2][2][2][2][3] = {{NTT_HI_DIR(false, false), NTT_HI_DIR(true, false)},
                                                 {NTT_HI_DIR(false, true), NTT_HI_DIR(true, true)}};
    return tab[full_tw ? 1 : 0][inverse ? 1 : 0][src_bitrev ? 1 : 0][has_pre ? 1 : 0][post];
}

// Pass LO: in place on rows of 2^b contiguous words; tile = 2^w_log rows.  grid (rows / W, n_cols)
template <bool INV, bool FULLTW>
__global__ void __launch_bounds__(NTT_THREADS, 3) ntt_pass_lo(NttKParams P) {
    extern __shared__ u64 sm_all[];
    const int b = P.b, w_log = P.w_log, pitch = P.pitch;
    const int S = 1 << b;
    u64 *tws = sm_all, *sm = sm_all + (FULLTW ? S : (S >> 1));
    const int n_el = 1 << (b + w_log);
    const size_t l0 = (size_t)blockIdx.x << (b + w_log);
    u64 *base = P.dst + (size_t)blockIdx.y * P.dst_stride + l0;
    load_twiddles(tws, P.wtab, b, FULLTW);
#if NTT_LO_VEC
    // 128-bit global accesses: a thread moves two adjacent points (they stay adjacent in the padded tile: the pad
    // words sit in front of multiples of 32 and the second point of a pair is odd)
    const bool vec = b >= 1 && n_el >= 2;
    if (vec) {
        const ulonglong2 *__restrict__ base2 = reinterpret_cast<const ulonglong2 *>(base);
        for (int q0 = threadIdx.x; q0 < n_el / 2; q0 += blockDim.x * NTT_IO_UNROLL) {
            ulonglong2 v[NTT_IO_UNROLL];
#pragma unroll
            for (int u = 0; u < NTT_IO_UNROLL; u++) {
                const int q = q0 + u * blockDim.x;
                if (q < n_el / 2) v[u] = __ldcs(base2 + q);
            }
#pragma unroll
            for (int u = 0; u < NTT_IO_UNROLL; u++) {
                const int e = 2 * (q0 + u * blockDim.x);
                if (e < n_el) {
                    u64 *d = sm + ntt_sm(e >> b, e & (S - 1), pitch);
                    d[0] = v[u].x;
                    d[1] = v[u].y;
                }
            }
        }
    } else
#endif
    for (int e0 = threadIdx.x; e0 < n_el; e0 += blockDim.x * NTT_IO_UNROLL) {
        u64 v[NTT_IO_UNROLL];
#pragma unroll
        for (int u = 0; u < NTT_IO_UNROLL; u++) {
            const int e = e0 + u * blockDim.x;
            if (e < n_el) v[u] = __ldcs(base + e);
        }
#pragma unroll
        for (int u = 0; u < NTT_IO_UNROLL; u++) {
            const int e = e0 + u * blockDim.x;
            if (e < n_el) sm[ntt_sm(e >> b, e & (S - 1), pitch)] = v[u];
        }
    }
    __syncthreads();
    dif_tile<INV, FULLTW>(sm, tws, b, w_log, pitch, P.groups);
    const bool scale = P.scale != 1 && !P.post_full;
#if NTT_LO_VEC
    if (vec && (P.n_dest == 0 || P.dest_lc >= 1)) {
        for (int q = threadIdx.x; q < n_el / 2; q += blockDim.x) {
            const int e = 2 * q;
            const u64 *srcp = sm + ntt_sm(e >> b, e & (S - 1), pitch);
            u64 v0 = srcp[0], v1 = srcp[1];
            if (scale) {
                v0 = gl_mul_c(v0, P.scale);
                v1 = gl_mul_c(v1, P.scale);
            }
            u64 *o = P.n_dest ? ntt_out(P, blockIdx.y, l0 + e) : base + e;
            __stcs(reinterpret_cast<ulonglong2 *>(o), make_ulonglong2(v0, v1));
        }
        return;
    }
#endif
    for (int e = threadIdx.x; e < n_el; e += blockDim.x) {
        u64 v = sm[ntt_sm(e >> b, e & (S - 1), pitch)];
        if (scale) v = gl_mul_c(v, P.scale);
        if (P.n_dest) __stcs(ntt_out(P, blockIdx.y, l0 + e), v);
        else __stcs(base + e, v);
    }
}

#if NTT_LO_TMA
// Pass LO with the tile moved by the copy engine (tiles of 2^7-point rows and up; ntt_pass_lo stays for the small ones and
// for a driver without cuTensorMapEncodeTiled): the tile is one contiguous run
// of 2^(b + w_log) words, described to the TMA unit as lines of 16 words; `cp.async.bulk.tensor.2d` brings it in as boxes
// of 256 lines with the 128-byte hardware swizzle (no pad words; dif_tile<.., XSW>), one thread arms the mbarrier and
// issues the copies, the others fill the twiddle table meanwhile.  The result goes back the same way -- into the local
// buffer, or into the slab of the rank that owns these leaves (NttDestMaps) -- unless it has to be scaled on the way.
__device__ __forceinline__ u32 ntt_smem_addr(const void *p) { return (u32)__cvta_generic_to_shared(p); }

// one tensor map per destination slab of a row-sharded output (NTT_LO_TMA_DEST): a tile of pass LO is 2^(b + w_log)
// consecutive leaves of one column, which lie inside ONE rank's slab, contiguously
struct NttDestMaps {
    CUtensorMap m[BSX_MAX_RANKS];
};
#ifndef NTT_LO_TMA_DEST
#define NTT_LO_TMA_DEST 1
#endif

template <bool INV, bool FULLTW>
__global__ void __launch_bounds__(NTT_THREADS, 3) ntt_pass_lo_tma(const __grid_constant__ CUtensorMap tmap,
                                                                  const __grid_constant__ NttDestMaps dmaps, NttKParams P) {
    extern __shared__ __align__(1024) u64 sm_tma[];
    const int b = P.b, w_log = P.w_log;
    const int S = 1 << b;
    const int n_el = 1 << (b + w_log);
    u64 *sm = sm_tma, *tws = sm_tma + n_el;
    u64 *bar = tws + (FULLTW ? S : (S >> 1));
    const size_t l0 = (size_t)blockIdx.x << (b + w_log);
    const size_t word0 = (size_t)blockIdx.y * P.dst_stride + l0;
    const int n_lines = n_el >> 4;
    const int box = n_lines < 256 ? n_lines : 256;
    const int line0 = (int)(word0 >> 4);
    const u32 bar_a = ntt_smem_addr(bar);
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_a) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar_a), "r"(n_el * 8) : "memory");
        for (int c = 0; c < n_lines; c += box)
            asm volatile(
                "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                ::"r"(ntt_smem_addr(sm + ((size_t)c << 4))), "l"(&tmap), "r"(0), "r"(line0 + c), "r"(bar_a)
                : "memory");
    }
    load_twiddles(tws, P.wtab, b, FULLTW);
    __syncthreads();  // the barrier word is initialised for everybody; the table is complete
    {
        u32 done = 0;
        while (!done)
            asm volatile(
                "{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}"
                : "=r"(done)
                : "r"(bar_a), "r"(0)
                : "memory");
    }
    dif_tile<INV, FULLTW, true>(sm, tws, b, w_log, S, P.groups);
    const bool scale = P.scale != 1 && !P.post_full;
    // row-sharded output: the tile goes to the slab of rank l0 >> dest_lc (peer memory over NVLink, or a local send
    // buffer), at column blockIdx.y of that slab; bulk store only when the tile lies inside one slab
    const bool dest_tma = NTT_LO_TMA_DEST && P.n_dest != 0 && P.dest_tma && P.dest_lc >= b + w_log;
    if (!scale && (P.n_dest == 0 || dest_tma)) {
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncthreads();
        if (threadIdx.x == 0) {
            const CUtensorMap *mp = &tmap;
            int line_out = line0;
            if (P.n_dest != 0) {
                mp = &dmaps.m[l0 >> P.dest_lc];
                line_out = (int)((((size_t)blockIdx.y << P.dest_lc) + (l0 & (((size_t)1 << P.dest_lc) - 1))) >> 4);
            }
            for (int c = 0; c < n_lines; c += box)
                asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.tile.bulk_group [%0, {%1, %2}], [%3];" ::"l"(mp),
                             "r"(0), "r"(line_out + c), "r"(ntt_smem_addr(sm + ((size_t)c << 4)))
                             : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            // local output: the tile only has to stay until it has been read; another rank's slab: until the writes are done
            if (P.n_dest != 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
            else asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        }
        return;
    }
    u64 *base = P.dst + word0;
    for (int q = threadIdx.x; q < n_el / 2; q += blockDim.x) {
        const int e = 2 * q;
        const u64 *srcp = sm + ((e >> b) << b) + ntt_xsw(e & (S - 1));
        u64 v0 = srcp[0], v1 = srcp[1];
        if (scale) {
            v0 = gl_mul_c(v0, P.scale);
            v1 = gl_mul_c(v1, P.scale);
        }
        if (P.n_dest && P.dest_lc < 1) {
            __stcs(ntt_out(P, blockIdx.y, l0 + e), v0);
            __stcs(ntt_out(P, blockIdx.y, l0 + e + 1), v1);
        } else {
            u64 *o = P.n_dest ? ntt_out(P, blockIdx.y, l0 + e) : base + e;
            __stcs(reinterpret_cast<ulonglong2 *>(o), make_ulonglong2(v0, v1));
        }
    }
}

typedef CUresult (*ntt_encode_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static ntt_encode_fn ntt_tensor_map_encoder() {
    static ntt_encode_fn fn = [] {
        void *f = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess)
            f = nullptr;
        return (ntt_encode_fn)f;
    }();
    return fn;
}
// the buffer `words` long at `base` as lines of 16 words, boxes of `box` lines, 128-byte swizzle
static bool ntt_lines_map(CUtensorMap *m, u64 *base, size_t words, int box) {
    ntt_encode_fn enc = ntt_tensor_map_encoder();
    if (!enc || ((uintptr_t)base & 127) || (words & 15) || (words >> 4) >= ((size_t)1 << 31)) return false;
    cuuint64_t dims[2] = {16, (cuuint64_t)(words >> 4)};
    cuuint64_t strides[1] = {128};
    cuuint32_t boxd[2] = {16, (cuuint32_t)box};
    cuuint32_t estr[2] = {1, 1};
    return enc(m, CU_TENSOR_MAP_DATA_TYPE_UINT64, 2, base, dims, strides, boxd, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
               CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}
#endif

// post_full[k][p][i0] = scale * pre_lo[k][i0] * omega_N^(i0 * bitrev_a(p)): everything pass HI multiplies
// into an output word besides the data, built once per (size, coset set) and cached in the context
__global__ void ntt_post_table_kernel(u64 *out, const u64 *pre_lo, const u64 *tw_lo, const u64 *tw_hi, u64 scale,
                                      int a, int b, int r_log) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >> (a + b + r_log)) return;
    u32 i0 = (u32)(i & (((size_t)1 << b) - 1));
    u32 p = (u32)((i >> b) & (((size_t)1 << a) - 1));
    u32 k = (u32)(i >> (a + b));
    u64 ex = (u64)i0 * bitrev32(p, a);
    u64 t = tw_lo[ex & 4095];
    if (ex >> 12) t = gl_mul(t, tw_hi[ex >> 12]);
    if (pre_lo) t = gl_mul(t, pre_lo[((size_t)k << b) + i0]);
    if (scale != 1) t = gl_mul(t, scale);
    out[i] = t;
}
void launch_ntt_post_table(u64 *out, const u64 *pre_lo, PowTable tw, u64 scale, int a, int b, int r_log,
                           cudaStream_t st) {
    size_t n = (size_t)1 << (a + b + r_log);
    BSX_COUNT_LAUNCH();
    ntt_post_table_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(out, pre_lo, tw.lo, tw.hi, scale, a, b, r_log);
}

int ntt_max_single_pass_log() { return NTT_MAX_A; }

// per-device opt-in to > 48 KB dynamic shared memory; called from bsx_ctx_create
void ntt_init_device() {
    for (int i = 0; i < 48; i++)
        cudaFuncSetAttribute(ntt_hi_variant(i & 1, (i >> 1) & 1, (i >> 2) % 3, (i / 12) & 1, i >= 24),
                             cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    cudaFuncSetAttribute(ntt_pass_lo<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    cudaFuncSetAttribute(ntt_pass_lo<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    cudaFuncSetAttribute(ntt_pass_lo<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    cudaFuncSetAttribute(ntt_pass_lo<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
#if NTT_LO_TMA
    cudaFuncSetAttribute(ntt_pass_lo_tma<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    cudaFuncSetAttribute(ntt_pass_lo_tma<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    cudaFuncSetAttribute(ntt_pass_lo_tma<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    cudaFuncSetAttribute(ntt_pass_lo_tma<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
#endif
}

// stage groups for a 2^bits-point tile transform: the last group (compile-time twiddles) takes up to
// four stages, the rest is split evenly into groups of at most four
static NttGroups plan_groups(int bits) {
    NttGroups G;
    G.n = 0;
    for (int i = 0; i < 4; i++) G.rho[i] = 0;
    if (bits <= 0) return G;
    int last = bits < 4 ? bits : 4;
    int rest = bits - last;
    int g = (rest + 3) / 4;
    for (int i = 0; i < g; i++) G.rho[G.n++] = rest / g + (i < rest % g ? 1 : 0);
    G.rho[G.n++] = last;
    return G;
}

// column pitch: 2^bits points + one pad word per 32, rounded up to 2 (mod 16)
static int ntt_pitch(int bits) {
    int p = (1 << bits) + ((1 << bits) >> 5) + ((1 << bits) >> 9);
    while ((p & 15) != 2) p++;
    return p;
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
    P.post_full = plan.b > 0 ? plan.post_full : nullptr;
    P.n_dest = plan.n_dest;
    P.dest_tma = 0;
    P.dest_lc = plan.dest_lc;
    for (int i = 0; i < BSX_MAX_RANKS; i++) P.dest[i] = i < plan.n_dest ? plan.dest[i] : nullptr;
    // pass HI: tile of 2^a points x W adjacent i0
    {
        int w_log = 13 - plan.a;  // 8192 elements per tile
        // 2^12-point tiles: two columns per tile (16-byte row pieces, completed in L2 by the neighbouring CTA)
        // keep the tile at 66 KB and three CTAs per SM; four columns would leave one 8-warp CTA per SM
        const int w_min = plan.a >= 12 ? NTT_W_MIN_A12 : 2;
        if (w_log < w_min) w_log = w_min;
        if (w_log > plan.b) w_log = plan.b;
        P.w_log = w_log;
        P.groups = plan_groups(plan.a);
        dim3 grid((unsigned)n_cols, (unsigned)(((size_t)1 << plan.b) >> w_log));
        size_t n_el = (size_t)1 << (plan.a + w_log);
        P.pitch = ntt_pitch(plan.a);
        size_t smem = sizeof(u64) * (((size_t)P.pitch << w_log) + ((size_t)1 << plan.a) / (ntt_tw_full(plan.a) ? 1 : 2));
        int threads = NTT_THREADS;
        while (threads > 32 && threads > (int)(n_el >> 1)) threads >>= 1;
        const int post = plan.b == 0 ? 0 : (P.post_full ? 2 : 1);
        BSX_COUNT_LAUNCH(); ntt_hi_variant(plan.src_bitrev != 0, plan.pre_hi != nullptr, post, plan.inverse != 0, ntt_tw_full(plan.a))<<<grid, threads, smem, st>>>(P);
    }
    if (plan.b > 0) {
        size_t rows = ((size_t)plan.n_cosets << plan.log_n) >> plan.b;  // per column
        int w_log = 13 - plan.b;
        if (w_log < 0) w_log = 0;
        while (((size_t)1 << w_log) > rows) w_log--;
        P.w_log = w_log;
        P.groups = plan_groups(plan.b);
        dim3 grid((unsigned)(rows >> w_log), (unsigned)n_cols);
        size_t n_el = (size_t)1 << (plan.b + w_log);
        P.pitch = ntt_pitch(plan.b);
        size_t smem = sizeof(u64) * (((size_t)P.pitch << w_log) + ((size_t)1 << plan.b) / (ntt_tw_full(plan.b) ? 1 : 2));
        int threads = NTT_THREADS;
        while (threads > 32 && threads > (int)(n_el >> 1)) threads >>= 1;
        BSX_COUNT_LAUNCH();
        const bool full = ntt_tw_full(plan.b);
#if NTT_LO_TMA
        CUtensorMap tmap;
        const size_t span = (size_t)(n_cols - 1) * dst_stride + ((size_t)plan.n_cosets << plan.log_n);
        const int box = (int)((n_el >> 4) < 256 ? (n_el >> 4) : 256);
        if (plan.b >= 7 && threads == NTT_THREADS && (dst_stride & 15) == 0 && ntt_lines_map(&tmap, dst, span, box)) {
            size_t smem_t = sizeof(u64) * (n_el + ((size_t)1 << plan.b) / (full ? 1 : 2) + 2);
            static NttDestMaps dmaps;  // only read by the kernel when P.dest_tma is set; copied into the launch
            NttDestMaps dm = dmaps;
            P.dest_tma = 0;
#if NTT_LO_TMA_DEST
            if (plan.n_dest > 0 && plan.dest_lc >= plan.b + w_log) {
                bool ok = true;  // slab of rank d: n_cols columns of 2^dest_lc leaves
                for (int d = 0; d < plan.n_dest && ok; d++) ok = ntt_lines_map(&dm.m[d], plan.dest[d], (size_t)n_cols << plan.dest_lc, box);
                P.dest_tma = ok ? 1 : 0;
            }
#endif
            if (plan.inverse) {
                if (full) ntt_pass_lo_tma<true, true><<<grid, threads, smem_t, st>>>(tmap, dm, P);
                else ntt_pass_lo_tma<true, false><<<grid, threads, smem_t, st>>>(tmap, dm, P);
            } else {
                if (full) ntt_pass_lo_tma<false, true><<<grid, threads, smem_t, st>>>(tmap, dm, P);
                else ntt_pass_lo_tma<false, false><<<grid, threads, smem_t, st>>>(tmap, dm, P);
            }
            return;
        }
#endif
        if (plan.inverse) {
            if (full) ntt_pass_lo<true, true><<<grid, threads, smem, st>>>(P);
            else ntt_pass_lo<true, false><<<grid, threads, smem, st>>>(P);
        } else {
            if (full) ntt_pass_lo<false, true><<<grid, threads, smem, st>>>(P);
            else ntt_pass_lo<false, false><<<grid, threads, smem, st>>>(P);
        }
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

// data[c][i] *= hi[i >> b] * lo[i & (2^b - 1)]: the coefficient scaling of PolynomialValues::coset_ifft
// (c_i *= shift^-i) from the two-level power tables of bsx_coset_tables
__global__ void scale_two_level_kernel(u64 *__restrict__ data, size_t stride, int log_n, int b,
                                       const u64 *__restrict__ hi, const u64 *__restrict__ lo) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >> log_n) return;
    u64 f = gl_mul_lazy(__ldg(hi + (i >> b)), __ldg(lo + (i & (((size_t)1 << b) - 1))));
    u64 *q = data + (size_t)blockIdx.y * stride + i;
    *q = gl_canon_fast(gl_mul_lazy(*q, f));
}
void launch_scale_two_level(u64 *data, size_t stride, int log_n, int b, int n_cols, const u64 *hi, const u64 *lo,
                            cudaStream_t st) {
    if (n_cols <= 0) return;
    size_t n = (size_t)1 << log_n;
    dim3 grid((unsigned)((n + 255) / 256), (unsigned)n_cols);
    BSX_COUNT_LAUNCH(); scale_two_level_kernel<<<grid, 256, 0, st>>>(data, stride, log_n, b, hi, lo);
}

// *flag |= 1 when any word of data[c][begin .. end) is non-zero (PolynomialCoeffs::trim_to_len's check)
__global__ void any_nonzero_kernel(const u64 *__restrict__ data, size_t stride, size_t begin, size_t end, int *flag) {
    size_t i = begin + (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < end && data[(size_t)blockIdx.y * stride + i] != 0) atomicOr(flag, 1);
}
void launch_any_nonzero(const u64 *data, size_t stride, size_t begin, size_t end, int n_cols, int *flag, cudaStream_t st) {
    if (n_cols <= 0 || end <= begin) return;
    dim3 grid((unsigned)((end - begin + 255) / 256), (unsigned)n_cols);
    BSX_COUNT_LAUNCH(); any_nonzero_kernel<<<grid, 256, 0, st>>>(data, stride, begin, end, flag);
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

// Salt columns when the caller supplies none (fri/oracle.rs lde_values: `F::rand_vec`, a CSPRNG upstream).
// ChaCha20 (RFC 8439 block function) in counter mode under a 256-bit key the host draws from the kernel's
// entropy pool per batch (api_batch.cu).  Block t yields the eight 64-bit candidates of outputs 8t .. 8t+7;
// a candidate >= p is rejected and replaced from that output's own spare stream (block counters with the
// top bit set), so the words are uniform in [0, p) and no opened salt word reveals anything about another.
__device__ __forceinline__ void chacha20_block(const u32 key[8], u64 counter, u32 out[16]) {
    u32 x[16] = {0x61707865u, 0x3320646eu, 0x79622d32u, 0x6b206574u, key[0], key[1], key[2], key[3],
                 key[4], key[5], key[6], key[7], (u32)counter, (u32)(counter >> 32), 0u, 0u};
    u32 s[16];
#pragma unroll
    for (int i = 0; i < 16; i++) s[i] = x[i];
#define CHACHA_QR(a, b, c, d)                               \
    x[a] += x[b]; x[d] = __funnelshift_l(x[d] ^ x[a], x[d] ^ x[a], 16); \
    x[c] += x[d]; x[b] = __funnelshift_l(x[b] ^ x[c], x[b] ^ x[c], 12); \
    x[a] += x[b]; x[d] = __funnelshift_l(x[d] ^ x[a], x[d] ^ x[a], 8);  \
    x[c] += x[d]; x[b] = __funnelshift_l(x[b] ^ x[c], x[b] ^ x[c], 7);
#pragma unroll 1
    for (int r = 0; r < 10; r++) {
        CHACHA_QR(0, 4, 8, 12) CHACHA_QR(1, 5, 9, 13) CHACHA_QR(2, 6, 10, 14) CHACHA_QR(3, 7, 11, 15)
        CHACHA_QR(0, 5, 10, 15) CHACHA_QR(1, 6, 11, 12) CHACHA_QR(2, 7, 8, 13) CHACHA_QR(3, 4, 9, 14)
    }
#undef CHACHA_QR
#pragma unroll
    for (int i = 0; i < 16; i++) out[i] = x[i] + s[i];
}
struct ChaChaKey {
    u32 w[8];
};
__global__ void fill_random_kernel(u64 *out, size_t n, ChaChaKey key, int raw) {
    size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t * 8 >= n) return;
    u32 blk[16];
    chacha20_block(key.w, (u64)t, blk);
#pragma unroll 1
    for (int j = 0; j < 8; j++) {
        const size_t i = t * 8 + j;
        if (i >= n) break;
        u64 v = ((u64)blk[2 * j + 1] << 32) | blk[2 * j];
        if (!raw) {
            u64 spare = 0;
            while (v >= GL_P) {  // probability 2^-32 per draw
                u32 ex[16];
                chacha20_block(key.w, (1ULL << 63) | ((u64)i << 8) | (spare >> 3), ex);
                v = ((u64)ex[2 * (spare & 7) + 1] << 32) | ex[2 * (spare & 7)];
                spare++;
            }
        }
        out[i] = v;
    }
}
// key: 8 words.  raw != 0: the key stream itself (test hook: RFC 8439 vectors), no rejection.
void launch_fill_random(u64 *out, size_t n, const u32 key[8], int raw, cudaStream_t st) {
    if (n == 0) return;
    ChaChaKey k;
    for (int i = 0; i < 8; i++) k.w[i] = key[i];
    const size_t blocks = (n + 7) / 8;
    BSX_COUNT_LAUNCH(); fill_random_kernel<<<(unsigned)((blocks + 127) / 128), 128, 0, st>>>(out, n, k, raw);
}
