// api_zs.cu -- the permutation argument's Z and partial-product polynomials on the device.
// Replaces plonky2 plonk/prover.rs `all_wires_permutation_partial_products` /
// `wires_permutation_partial_products_and_zs` and plonk/vanishing_poly.rs-independent helpers
// `quotient_chunk_products`, `partial_products_and_z_gx` (plonk/permutation_argument / util/partial_products.rs):
// the prover step between the wires commitment and the Zs commitment.  Gate independent: it reads the routed wire
// values, the sigma values, k_is and the challenges only.
//
// Per row i (x = omega^i) and challenge (beta, gamma), chunk m of `max_degree` routed wires:
//     q_m = prod_j (w_j + beta k_j x + gamma) / prod_j (w_j + beta sigma_j + gamma)
// Z(omega^0) = 1, Z(g x) = Z(x) prod_m q_m, partial product k of the row = Z(x) q_0 ... q_k.
// Field arithmetic is exact, so any evaluation order gives plonky2's words: one thread per (row, challenge) forms the
// chunk numerators / denominators and inverts the denominators together (one inversion per row), a multiplicative
// scan over the rows gives Z, a last pass scales the running products by Z(x).
#include "api_internal.h"

#define ZS_MAX_CHUNKS 32

__global__ void __launch_bounds__(256) zs_rows_kernel(const u64 *__restrict__ wires, const u64 *__restrict__ sigmas,
                                                      size_t n, int n_routed, int max_degree, const u64 *__restrict__ k_is,
                                                      const u64 *__restrict__ betas, const u64 *__restrict__ gammas,
                                                      const u64 *__restrict__ xs, int n_chunks, u64 *__restrict__ running,
                                                      u64 *__restrict__ row_prod, int *__restrict__ zero_flag) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int c = blockIdx.y;
    const u64 beta = betas[c], gamma = gammas[c];
    const u64 bx = gl_mul(beta, xs[i]);
    // lazy arithmetic (any u64 representative) inside the row, canonical words at the stores
    u64 num[ZS_MAX_CHUNKS], den[ZS_MAX_CHUNKS];
    for (int m = 0; m < n_chunks; m++) {
        u64 a = 1, d = 1;
        const int j1 = min(n_routed, (m + 1) * max_degree);
        for (int j = m * max_degree; j < j1; j++) {
            u64 wg = gl_add_lazy_canon(wires[(size_t)j * n + i], gamma);            // w + gamma
            a = gl_mul_lazy(a, gl_mul_add_lazy(bx, k_is[j], wg));                   // . (w + beta k x + gamma)
            d = gl_mul_lazy(d, gl_mul_add_lazy(beta, sigmas[(size_t)j * n + i], wg));  // . (w + beta sigma + gamma)
        }
        num[m] = a;
        den[m] = d;
    }
    // batch inversion of the chunk denominators; a zero denominator is where upstream's
    // batch_multiplicative_inverse panics ("Tried to invert zero"): flagged, the call fails
    u64 pre[ZS_MAX_CHUNKS];
    u64 acc = 1;
    for (int m = 0; m < n_chunks; m++) {
        pre[m] = acc;
        acc = gl_mul_lazy(acc, den[m]);
    }
    if (gl_canon_fast(acc) == 0) atomicOr(zero_flag, 1);
    u64 inv = 1, sq = acc;  // acc^(p - 2), p - 2 = 0xFFFFFFFEFFFFFFFF
    for (int bit = 0; bit < 64; bit++) {
        if (bit != 32) inv = gl_mul_lazy(inv, sq);
        sq = gl_mul_lazy(sq, sq);
    }
    for (int m = n_chunks - 1; m >= 0; m--) {
        u64 dinv = gl_mul_lazy(inv, pre[m]);
        inv = gl_mul_lazy(inv, den[m]);
        num[m] = gl_mul_lazy(num[m], dinv);  // q_m
    }
    acc = 1;
    for (int m = 0; m < n_chunks; m++) {
        acc = gl_mul_lazy(acc, num[m]);
        if (m + 1 < n_chunks) running[((size_t)c * (n_chunks - 1) + m) * n + i] = gl_canon_fast(acc);
    }
    acc = gl_canon_fast(acc);
    row_prod[(size_t)c * n + i] = acc;
}

// exclusive multiplicative scan, three launches: per-block products, scan of the block products, apply
#define ZS_SCAN_ITEMS 4
#define ZS_SCAN_BLOCK (256 * ZS_SCAN_ITEMS)
__device__ __forceinline__ u64 zs_block_scan(u64 v, u64 *sh) {  // inclusive product over the 256 threads
    sh[threadIdx.x] = v;
    __syncthreads();
    for (int off = 1; off < 256; off <<= 1) {
        u64 t = threadIdx.x >= off ? sh[threadIdx.x - off] : 1;
        __syncthreads();
        sh[threadIdx.x] = v = gl_mul(v, t);
        __syncthreads();
    }
    return v;
}
__global__ void __launch_bounds__(256) zs_block_prod_kernel(const u64 *__restrict__ q, size_t n, u64 *__restrict__ block_prod) {
    __shared__ u64 sh[256];
    const u64 *col = q + (size_t)blockIdx.y * n;
    size_t base = (size_t)blockIdx.x * ZS_SCAN_BLOCK + (size_t)threadIdx.x * ZS_SCAN_ITEMS;
    u64 v = 1;
    for (int t = 0; t < ZS_SCAN_ITEMS; t++)
        if (base + t < n) v = gl_mul(v, col[base + t]);
    v = zs_block_scan(v, sh);
    if (threadIdx.x == 255) block_prod[(size_t)blockIdx.y * gridDim.x + blockIdx.x] = v;
}
__global__ void __launch_bounds__(256) zs_scan_blocks_kernel(u64 *block_prod, int n_blocks) {  // in place -> exclusive
    __shared__ u64 sh[256];
    u64 *col = block_prod + (size_t)blockIdx.x * n_blocks;
    u64 carry = 1;
    for (int b0 = 0; b0 < n_blocks; b0 += 256) {
        int b = b0 + threadIdx.x;
        u64 v = b < n_blocks ? col[b] : 1;
        u64 inc = zs_block_scan(v, sh);
        u64 prev = threadIdx.x ? sh[threadIdx.x - 1] : 1;
        u64 total = sh[255];
        __syncthreads();
        if (b < n_blocks) col[b] = gl_mul(carry, prev);
        carry = gl_mul(carry, total);
        (void)inc;
    }
}
// z[i] = prod_{t < i} q[t]; out_z column c, and the partial products of column c scaled by z
__global__ void __launch_bounds__(256) zs_apply_kernel(const u64 *__restrict__ q, size_t n, const u64 *__restrict__ block_excl,
                                                       int n_chall, int n_pp, u64 *__restrict__ out) {
    __shared__ u64 sh[256];
    const int c = blockIdx.y;
    const u64 *col = q + (size_t)c * n;
    size_t base = (size_t)blockIdx.x * ZS_SCAN_BLOCK + (size_t)threadIdx.x * ZS_SCAN_ITEMS;
    u64 loc[ZS_SCAN_ITEMS], v = 1;
    for (int t = 0; t < ZS_SCAN_ITEMS; t++) {
        loc[t] = base + t < n ? col[base + t] : 1;
        v = gl_mul(v, loc[t]);
    }
    zs_block_scan(v, sh);
    u64 z = gl_mul(block_excl[(size_t)c * gridDim.x + blockIdx.x], threadIdx.x ? sh[threadIdx.x - 1] : 1);
    for (int t = 0; t < ZS_SCAN_ITEMS; t++) {
        size_t i = base + t;
        if (i >= n) break;
        out[(size_t)c * n + i] = z;  // Zs first (plonk/prover.rs: "Z is expected at the front of our batch")
        for (int m = 0; m < n_pp; m++) {
            u64 *p = out + ((size_t)n_chall + (size_t)c * n_pp + m) * n + i;
            *p = gl_mul(*p, z);
        }
        z = gl_mul(z, loc[t]);
    }
}

extern "C" int32_t bsx_partial_products_and_zs(bsx_ctx *ctx, const uint64_t *wires, const uint64_t *sigmas, int32_t mem,
                                               int32_t log_n, int32_t n_routed, int32_t max_degree,
                                               const uint64_t *k_is_host, const uint64_t *betas_host,
                                               const uint64_t *gammas_host, int32_t n_challenges, uint64_t *out,
                                               int32_t out_mem) {
    if (!ctx || !wires || !sigmas || !k_is_host || !betas_host || !gammas_host || !out) return BSX_ERR_INVALID_ARG;
    BsxScope scope(ctx);
    if (log_n < 0 || log_n > 24) return bsx_fail(ctx, BSX_ERR_UNSUPPORTED, "log_n=%d outside [0, 24]", log_n);
    if (max_degree < 2) return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "max_degree=%d: quotient_chunk_products needs > 1", max_degree);
    if (n_routed < 1 || n_challenges < 1 || n_challenges > 64) return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "bad n_routed/n_challenges");
    const int n_chunks = (n_routed + max_degree - 1) / max_degree, n_pp = n_chunks - 1;
    if (n_chunks > ZS_MAX_CHUNKS)
        return bsx_fail(ctx, BSX_ERR_UNSUPPORTED, "%d chunks of %d routed wires (at most %d)", n_chunks, max_degree, ZS_MAX_CHUNKS);
    CU(ctx, cudaSetDevice(ctx->device));
    const size_t n = (size_t)1 << log_n, in_words = n * (size_t)n_routed, out_words = n * (size_t)n_challenges * n_chunks;
    const int n_blocks = (int)((n + ZS_SCAN_BLOCK - 1) / ZS_SCAN_BLOCK);
    u64 *w_dev = nullptr, *s_dev = nullptr, *consts, *xs, *res, *q, *bp, *flag;
    if (mem == BSX_HOST) {
        TRY(bsx_alloc(ctx, &w_dev, in_words));
        TRY(bsx_alloc(ctx, &s_dev, in_words));
        TRY(bsx_copy_in(ctx, w_dev, (const u64 *)wires, in_words, mem));
        TRY(bsx_copy_in(ctx, s_dev, (const u64 *)sigmas, in_words, mem));
    }
    const u64 *w = mem == BSX_HOST ? w_dev : (const u64 *)wires, *s = mem == BSX_HOST ? s_dev : (const u64 *)sigmas;
    TRY(bsx_alloc(ctx, &consts, (size_t)n_routed + 2 * n_challenges));
    TRY(bsx_alloc(ctx, &xs, n));
    TRY(bsx_alloc(ctx, &q, n * (size_t)n_challenges));
    TRY(bsx_alloc(ctx, &bp, (size_t)n_blocks * n_challenges));
    TRY(bsx_alloc(ctx, &flag, 1));
    CU(ctx, cudaMemsetAsync(flag, 0, sizeof(u64), ctx->stream));
    if (out_mem == BSX_HOST) TRY(bsx_alloc(ctx, &res, out_words));
    else res = (u64 *)out;
    std::vector<u64> hc((size_t)n_routed + 2 * n_challenges);
    for (int j = 0; j < n_routed; j++) hc[j] = k_is_host[j] % GL_P;
    for (int c = 0; c < n_challenges; c++) {
        hc[n_routed + c] = betas_host[c] % GL_P;
        hc[n_routed + n_challenges + c] = gammas_host[c] % GL_P;
    }
    TRY(bsx_copy_in(ctx, consts, hc.data(), hc.size(), BSX_HOST));
    launch_fill_powers(xs, gl_host_root(log_n), 1, n, ctx->stream);  // F::two_adic_subgroup(degree_bits)
    dim3 grid_rows((unsigned)((n + 255) / 256), (unsigned)n_challenges);
    BSX_COUNT_LAUNCH(); zs_rows_kernel<<<grid_rows, 256, 0, ctx->stream>>>(w, s, n, n_routed, max_degree, consts, consts + n_routed,
                                                      consts + n_routed + n_challenges, xs, n_chunks,
                                                      res + n * (size_t)n_challenges, q, (int *)flag);
    dim3 grid_scan((unsigned)n_blocks, (unsigned)n_challenges);
    BSX_COUNT_LAUNCH(); zs_block_prod_kernel<<<grid_scan, 256, 0, ctx->stream>>>(q, n, bp);
    BSX_COUNT_LAUNCH(); zs_scan_blocks_kernel<<<n_challenges, 256, 0, ctx->stream>>>(bp, n_blocks);
    BSX_COUNT_LAUNCH(); zs_apply_kernel<<<grid_scan, 256, 0, ctx->stream>>>(q, n, bp, n_challenges, n_pp, res);
    CU(ctx, cudaGetLastError());
    int host_flag = 0;
    CU(ctx, cudaMemcpyAsync(&host_flag, flag, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    if (out_mem == BSX_HOST) TRY(bsx_copy_out(ctx, (u64 *)out, res, out_words, BSX_HOST));
    CU(ctx, cudaStreamSynchronize(ctx->stream));
    if (host_flag)
        return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "zero denominator in the permutation quotients (upstream panics: tried to invert zero)");
    return BSX_OK;
}
