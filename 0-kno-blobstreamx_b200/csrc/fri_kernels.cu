// fri_kernels.cu -- device side of the FRI half of the path (K6, K8 and the challenger).
// Replaces plonky2 iop/challenger.rs (Challenger), fri/oracle.rs `prove_openings` (alpha-combine,
// divide_by_linear), fri/prover.rs (`fri_committed_trees` fold, `fri_proof_of_work`,
// `fri_prover_query_rounds`).  Extension elements F_p[x]/(x^2-7) are kept PLANAR on device
// (plane 0 = c0, plane 1 = c1) so both planes go through the base-field NTT unchanged.
#include "internal.h"
#include "poseidon.cuh"

// ------------------------------------------------------------------ challenger (one warp)
// Layout identical to include/bsx_b200.h bsx_challenger.
struct DevChallenger {
    u64 state[12];
    u64 in[8];
    u64 out[8];
    int n_in, n_out;
};

// every function below is executed by all 32 lanes of the kernel's single warp on a challenger in shared memory
__device__ void chal_duplex(DevChallenger *c, int lane) {
    __syncwarp();
    u64 x = lane < 12 ? c->state[lane] : 0;
    if (lane < c->n_in) x = c->in[lane];  // overwrite mode
    x = poseidon_permute_warp(x, lane);
    __syncwarp();
    if (lane < 12) c->state[lane] = x;
    if (lane < 8) c->out[lane] = x;
    if (lane == 0) {
        c->n_in = 0;
        c->n_out = 8;
    }
    __syncwarp();
}
__device__ void chal_observe(DevChallenger *c, u64 x, int lane) {
    const int n_in = c->n_in;  // uniform
    __syncwarp();
    if (lane == 0) {
        c->n_out = 0;
        c->in[n_in] = x;
        c->n_in = n_in + 1;
    }
    __syncwarp();
    if (n_in + 1 == 8) chal_duplex(c, lane);
}
__device__ u64 chal_get(DevChallenger *c, int lane) {
    if (c->n_in != 0 || c->n_out == 0) chal_duplex(c, lane);
    const int n_out = c->n_out;  // uniform
    const u64 v = c->out[n_out - 1];
    __syncwarp();
    if (lane == 0) c->n_out = n_out - 1;
    __syncwarp();
    return v;
}
__device__ void chal_load(DevChallenger *sc, const DevChallenger *c, int lane) {
    const u64 *src = (const u64 *)c;
    u64 *dst = (u64 *)sc;
    for (int i = lane; i < (int)(sizeof(DevChallenger) / 8); i += 32) dst[i] = src[i];
    __syncwarp();
}
__device__ void chal_store(DevChallenger *c, const DevChallenger *sc, int lane) {
    __syncwarp();
    const u64 *src = (const u64 *)sc;
    u64 *dst = (u64 *)c;
    for (int i = lane; i < (int)(sizeof(DevChallenger) / 8); i += 32) dst[i] = src[i];
}

// mode 0: n contiguous words.  mode 1: n words of an interleaved extension stream read from planar
// storage: word i = plane (i & 1), index i >> 1.
__global__ void chal_observe_kernel(DevChallenger *c, const u64 *data, size_t n, int mode, size_t plane_stride) {
    __shared__ DevChallenger sc;
    const int lane = threadIdx.x;
    chal_load(&sc, c, lane);
    for (size_t i = 0; i < n; i++) chal_observe(&sc, mode ? data[(i & 1) * plane_stride + (i >> 1)] : data[i], lane);
    chal_store(c, &sc, lane);
}
__global__ void chal_get_kernel(DevChallenger *c, u64 *out, int n, u64 modulus) {
    __shared__ DevChallenger sc;
    const int lane = threadIdx.x;
    chal_load(&sc, c, lane);
    for (int i = 0; i < n; i++) {
        u64 v = chal_get(&sc, lane);
        if (lane == 0) out[i] = modulus ? v % modulus : v;
    }
    chal_store(c, &sc, lane);
}
void launch_chal_observe(void *c, const u64 *data, size_t n, int mode, size_t plane_stride, cudaStream_t st) {
    if (n == 0) return;
    BSX_COUNT_LAUNCH(); chal_observe_kernel<<<1, 32, 0, st>>>((DevChallenger *)c, data, n, mode, plane_stride);
}
void launch_chal_get(void *c, u64 *out, int n, u64 modulus, cudaStream_t st) {
    if (n <= 0) return;
    BSX_COUNT_LAUNCH(); chal_get_kernel<<<1, 32, 0, st>>>((DevChallenger *)c, out, n, modulus);
}

// ------------------------------------------------------------------ proof of work
// fri/prover.rs fri_proof_of_work: candidate w goes into lane n_in of the duplex state; the
// response is lane 7 of the permuted state.  Every candidate of [base, base + n) is tried and the
// SMALLEST valid one is kept (atomicMin), which makes the witness deterministic.
__global__ void __launch_bounds__(256) pow_kernel(const DevChallenger *c, u64 base, u64 n, int pow_bits, u64 *best) {
    u64 i = (u64)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    u64 cand = base + i;
    if (cand >= GL_P) return;
    u64 s[12];
#pragma unroll
    for (int t = 0; t < 12; t++) s[t] = c->state[t];
    const int n_in = c->n_in;
#pragma unroll
    for (int t = 0; t < 8; t++)
        if (t < n_in) s[t] = c->in[t];
#pragma unroll
    for (int t = 0; t < 8; t++)
        if (t == n_in) s[t] = cand;
    poseidon_permute_lazy(s);
    u64 resp = gl_canon(s[7]);
    if (__clzll((long long)resp) >= pow_bits) atomicMin((unsigned long long *)best, (unsigned long long)cand);
}
void launch_pow(const void *c, u64 base, u64 n, int pow_bits, u64 *best, cudaStream_t st) {
    BSX_COUNT_LAUNCH(); pow_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>((const DevChallenger *)c, base, n, pow_bits, best);
}
// observe the witness and draw the response (kept on device)
void launch_chal_pow_finish(void *c, const u64 *witness, u64 *response, cudaStream_t st) {
    launch_chal_observe(c, witness, 1, 0, 0, st);
    launch_chal_get(c, response, 1, 0, st);
}

// ------------------------------------------------------------------ prove_openings combine
// pw[j] = alpha^j, j < n (extension, interleaved), single thread -- n is a few hundred
__global__ void ext_powers_kernel(const u64 *alpha, u64 *pw, int n) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    gl2 a = {alpha[0], alpha[1]}, cur = {1, 0};
    for (int j = 0; j < n; j++) {
        pw[2 * j] = cur.c0;
        pw[2 * j + 1] = cur.c1;
        cur = gl2_mul(cur, a);
    }
}
void launch_ext_powers(const u64 *alpha, u64 *pw, int n, cudaStream_t st) {
    if (n <= 0) return;
    BSX_COUNT_LAUNCH(); ext_powers_kernel<<<1, 32, 0, st>>>(alpha, pw, n);
}

// comp[i] (+)= sum_j pw[j0 + j] * poly_j[i] for the polynomials listed in ptrs (device array of
// column pointers), positionwise in STORAGE order (the caller groups columns by storage order).
__global__ void __launch_bounds__(256) combine_kernel(const u64 *const *__restrict__ ptrs, const int *__restrict__ pw_index,
                                                      int n_polys, const u64 *__restrict__ pw, size_t n, u64 *comp0,
                                                      u64 *comp1, int accumulate) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    // lazy 160-bit accumulators: up to 2^12 products fit (x4 < 2^32)
    gl_acc160 A0 = {0, 0, 0, 0, 0}, A1 = {0, 0, 0, 0, 0};
    for (int j = 0; j < n_polys; j++) {
        u64 v = __ldg(ptrs[j] + i);
        int k = pw_index[j];
        gl_acc_mac(A0, v, pw[2 * k]);
        gl_acc_mac(A1, v, pw[2 * k + 1]);
    }
    u64 r0 = gl_canon(gl_acc_reduce(A0)), r1 = gl_canon(gl_acc_reduce(A1));
    if (accumulate) {
        r0 = gl_add(r0, comp0[i]);
        r1 = gl_add(r1, comp1[i]);
    }
    comp0[i] = r0;
    comp1[i] = r1;
}
void launch_combine(const u64 *const *ptrs, const int *pw_index, int n_polys, const u64 *pw, size_t n, u64 *comp0,
                    u64 *comp1, int accumulate, cudaStream_t st) {
    if (n == 0) return;
    BSX_COUNT_LAUNCH(); combine_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(ptrs, pw_index, n_polys, pw, n, comp0, comp1, accumulate);
}

// ------------------------------------------------------------------ divide_by_linear (suffix Horner scan)
// polynomial/mod.rs divide_by_linear: acc_i = acc_{i+1} * z + c_i from the top; quotient q_{i-1} = acc_i.
// Hierarchical: level L holds chunk sums with multiplier m_L = z^(16^L).
#define SCAN_CHUNK 16
__global__ void scan_mults_kernel(const u64 *z, u64 *m, int levels) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    gl2 cur = {z[0], z[1]};
    for (int l = 0; l < levels; l++) {
        m[2 * l] = cur.c0;
        m[2 * l + 1] = cur.c1;
        for (int k = 0; k < 4; k++) cur = gl2_mul(cur, cur);  // ^16
    }
}
__global__ void __launch_bounds__(256) scan_up_kernel(const u64 *in0, const u64 *in1, size_t n_in, const u64 *m,
                                                      u64 *out0, u64 *out1, size_t n_out) {
    size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_out) return;
    gl2 mm = {m[0], m[1]}, acc = {0, 0};
    for (int j = SCAN_CHUNK - 1; j >= 0; j--) {
        size_t i = t * SCAN_CHUNK + j;
        if (i < n_in) acc = gl2_add(gl2_mul(acc, mm), gl2{in0[i], in1[i]});
    }
    out0[t] = acc.c0;
    out1[t] = acc.c1;
}
// suffix at this level including the carry from the level above; at level 0 (fin != null) the
// result is folded into the X-multiplied final polynomial: F[i] = F[i] * sh + acc_i (i >= 1), F[0] = 0.
__global__ void __launch_bounds__(256) scan_down_kernel(const u64 *val0, const u64 *val1, size_t n, const u64 *m,
                                                        const u64 *up0, const u64 *up1, size_t n_up, u64 *suf0,
                                                        u64 *suf1, u64 *fin0, u64 *fin1, const u64 *sh) {
    size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t * SCAN_CHUNK >= n) return;
    gl2 mm = {m[0], m[1]};
    gl2 acc = {0, 0};
    if (up0 && t + 1 < n_up) acc = gl2{up0[t + 1], up1[t + 1]};
    gl2 s = {1, 0};
    if (fin0) s = gl2{sh[0], sh[1]};
    for (int j = SCAN_CHUNK - 1; j >= 0; j--) {
        size_t i = t * SCAN_CHUNK + j;
        if (i >= n) continue;
        acc = gl2_add(gl2_mul(acc, mm), gl2{val0[i], val1[i]});
        if (fin0) {
            gl2 f = {0, 0};
            if (i >= 1) f = gl2_add(gl2_mul(gl2{fin0[i], fin1[i]}, s), acc);
            fin0[i] = f.c0;
            fin1[i] = f.c1;
        } else {
            suf0[i] = acc.c0;
            suf1[i] = acc.c1;
        }
    }
}
// fin (planar, length n, already multiplied by X) <- fin * sh + X * ((comp(X) - comp(z)) / (X - z))
// work: 2 planes x (n/16 + n/256 + ... ) words x 2 (values and suffixes) + 2 * levels multipliers
size_t divide_by_linear_workspace_words(size_t n) {
    size_t w = 64, len = n;
    while (len > 1) {
        len = (len + SCAN_CHUNK - 1) / SCAN_CHUNK;
        w += 4 * len;
    }
    return w;
}
void launch_divide_by_linear_accumulate(const u64 *comp0, const u64 *comp1, size_t n, const u64 *z, const u64 *sh,
                                        u64 *fin0, u64 *fin1, u64 *work, cudaStream_t st) {
    // level sizes
    size_t lens[16];
    int levels = 0;
    lens[0] = n;
    while (lens[levels] > 1) {
        lens[levels + 1] = (lens[levels] + SCAN_CHUNK - 1) / SCAN_CHUNK;
        levels++;
    }
    u64 *m = work;  // 2 * (levels + 1) words
    BSX_COUNT_LAUNCH(); scan_mults_kernel<<<1, 32, 0, st>>>(z, m, levels + 1);
    u64 *p = work + 64;
    const u64 *v0[16], *v1[16];
    u64 *s0[16], *s1[16];
    v0[0] = comp0;
    v1[0] = comp1;
    for (int l = 1; l <= levels; l++) {
        u64 *a = p, *b = p + lens[l], *c = p + 2 * lens[l], *d = p + 3 * lens[l];
        p += 4 * lens[l];
        v0[l] = a; v1[l] = b; s0[l] = c; s1[l] = d;
        BSX_COUNT_LAUNCH(); scan_up_kernel<<<(unsigned)((lens[l] + 255) / 256), 256, 0, st>>>(v0[l - 1], v1[l - 1], lens[l - 1], m + 2 * (l - 1), a, b, lens[l]);
    }
    // downward: the top level has length 1 and no carry
    for (int l = levels; l >= 0; l--) {
        const u64 *u0 = l < levels ? s0[l + 1] : nullptr, *u1 = l < levels ? s1[l + 1] : nullptr;
        size_t n_up = l < levels ? lens[l + 1] : 0;
        size_t threads = (lens[l] + SCAN_CHUNK - 1) / SCAN_CHUNK;
        BSX_COUNT_LAUNCH(); scan_down_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, st>>>(
            v0[l], v1[l], lens[l], m + 2 * l, u0, u1, n_up, l ? s0[l] : nullptr, l ? s1[l] : nullptr,
            l ? nullptr : fin0, l ? nullptr : fin1, sh);
    }
}

// ------------------------------------------------------------------ FRI fold (K6)
// fri/prover.rs: coeffs <- chunks_exact(arity).map(|c| reduce_with_powers(c, beta)); planar in/out
__global__ void __launch_bounds__(256) fri_fold_kernel(const u64 *in0, const u64 *in1, size_t n_out, int arity_bits,
                                                       const u64 *beta, u64 *out0, u64 *out1) {
    size_t j = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n_out) return;
    gl2 b = {beta[0], beta[1]}, acc = {0, 0};
    const int arity = 1 << arity_bits;
    for (int k = arity - 1; k >= 0; k--) {
        size_t i = ((size_t)j << arity_bits) + k;
        acc = gl2_add(gl2_mul(acc, b), gl2{in0[i], in1[i]});
    }
    out0[j] = acc.c0;
    out1[j] = acc.c1;
}
void launch_fri_fold(const u64 *in0, const u64 *in1, size_t n_out, int arity_bits, const u64 *beta, u64 *out0,
                     u64 *out1, cudaStream_t st) {
    if (n_out == 0) return;
    BSX_COUNT_LAUNCH(); fri_fold_kernel<<<(unsigned)((n_out + 255) / 256), 256, 0, st>>>(in0, in1, n_out, arity_bits, beta, out0, out1);
}

// values (planar, already in bit-reversed order) -> column-major leaf matrix of the round's tree:
// leaf j = [v[arity*j + t] for t < arity] flattened (c0, c1): column 2t+e, row j
__global__ void __launch_bounds__(256) fri_leaves_kernel(const u64 *v0, const u64 *v1, size_t len, int arity_bits,
                                                         u64 *leaves) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= len) return;
    size_t j = i >> arity_bits, t = i & (((size_t)1 << arity_bits) - 1), nl = len >> arity_bits;
    leaves[(2 * t) * nl + j] = v0[i];
    leaves[(2 * t + 1) * nl + j] = v1[i];
}
void launch_fri_leaves(const u64 *v0, const u64 *v1, size_t len, int arity_bits, u64 *leaves, cudaStream_t st) {
    if (len == 0) return;
    BSX_COUNT_LAUNCH(); fri_leaves_kernel<<<(unsigned)((len + 255) / 256), 256, 0, st>>>(v0, v1, len, arity_bits, leaves);
}

// ------------------------------------------------------------------ query openings (K8)
// For query q: x = idx[q] >> shift.  Writes leaf row (width words) then sibling digests bottom-up
// into out[q * q_stride + off ...].
__global__ void fri_query_gather_kernel(const u64 *idx, int shift, const u64 *leaves, size_t col_stride, int width,
                                        const u64 *digests, int log_leaves, int n_sib, u64 *out, size_t q_stride,
                                        size_t off) {
    int q = blockIdx.x;
    size_t x = (size_t)(idx[q] >> shift);
    u64 *o = out + (size_t)q * q_stride + off;
    for (int c = threadIdx.x; c < width; c += blockDim.x) o[c] = leaves[(size_t)c * col_stride + x];
    for (int e = threadIdx.x; e < 4 * n_sib; e += blockDim.x) {
        int k = e >> 2, w = e & 3;
        size_t lvl_off = 0;
        for (int j = 0; j < k; j++) lvl_off += (size_t)1 << (log_leaves - j);
        o[width + e] = digests[4 * (lvl_off + ((x >> k) ^ 1)) + w];
    }
}
void launch_fri_query_gather(const u64 *idx, int n_queries, int shift, const u64 *leaves, size_t col_stride, int width,
                             const u64 *digests, int log_leaves, int n_sib, u64 *out, size_t q_stride, size_t off,
                             cudaStream_t st) {
    if (n_queries <= 0) return;
    BSX_COUNT_LAUNCH(); fri_query_gather_kernel<<<n_queries, 128, 0, st>>>(idx, shift, leaves, col_stride, width, digests, log_leaves, n_sib, out, q_stride, off);
}

// planar extension (2 planes, stride plane_stride) -> interleaved [c0, c1, ...], n elements
__global__ void ext_interleave_kernel(const u64 *planar, size_t plane_stride, size_t n, u64 *out) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    out[2 * i] = planar[i];
    out[2 * i + 1] = planar[plane_stride + i];
}
void launch_ext_interleave(const u64 *planar, size_t plane_stride, size_t n, u64 *out, cudaStream_t st) {
    if (n == 0) return;
    BSX_COUNT_LAUNCH(); ext_interleave_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(planar, plane_stride, n, out);
}
__global__ void ext_deinterleave_kernel(const u64 *in, size_t n, u64 *planar, size_t plane_stride) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    planar[i] = in[2 * i];
    planar[plane_stride + i] = in[2 * i + 1];
}
void launch_ext_deinterleave(const u64 *in, size_t n, u64 *planar, size_t plane_stride, cudaStream_t st) {
    if (n == 0) return;
    BSX_COUNT_LAUNCH(); ext_deinterleave_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(in, n, planar, plane_stride);
}
// sh = alpha^k (extension), single thread
__global__ void ext_pow_kernel(const u64 *a, u64 k, u64 *out) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    gl2 r = gl2_pow(gl2{a[0], a[1]}, k);
    out[0] = r.c0;
    out[1] = r.c1;
}
void launch_ext_pow(const u64 *a, u64 k, u64 *out, cudaStream_t st) {
    BSX_COUNT_LAUNCH(); ext_pow_kernel<<<1, 32, 0, st>>>(a, k, out);
}

// ------------------------------------------------------------------ sum of k vectors (mod p)
__global__ void __launch_bounds__(256) field_sum_kernel(const u64 *__restrict__ parts, int k, size_t words, u64 *__restrict__ out) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= words) return;
    u64 acc = 0;
    for (int j = 0; j < k; j++) acc = gl_add(acc, gl_canon(parts[(size_t)j * words + i]));
    out[i] = acc;
}
void launch_field_sum(const u64 *parts, int k, size_t words, u64 *out, cudaStream_t st) {
    if (words == 0) return;
    BSX_COUNT_LAUNCH(); field_sum_kernel<<<(unsigned)((words + 255) / 256), 256, 0, st>>>(parts, k, words, out);
}
