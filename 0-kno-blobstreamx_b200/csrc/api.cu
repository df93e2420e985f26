// api.cu -- the C ABI of libbsx_b200.so (include/bsx_b200.h): contexts, device-resident
// PolynomialBatch / MerkleTree handles and their accessors.  Host orchestration only; all field
// arithmetic happens in the kernels of ntt_kernels.cu / hash_kernels.cu / fri_kernels.cu.
// There is deliberately no CPU fallback anywhere in this file.
#include "api_internal.h"

#include <cstdarg>
#include <cstdio>
#include <cstring>

// ------------------------------------------------------------------------------------ helpers
int bsx_fail(bsx_ctx *ctx, int code, const char *fmt, ...) {
    if (ctx) {
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(ctx->err, sizeof(ctx->err), fmt, ap);
        va_end(ap);
        if (getenv("BSX_LOG")) fprintf(stderr, "[bsx_b200] error %d: %s\n", code, ctx->err);
    }
    return code;
}

// Buffers of BSX_BIG_BLOCK_BYTES and more (coefficients, LDEs, digest levels, transform scratch) are recycled by
// EXACT size through a per-context cache instead of going back to the stream-ordered pool: a prover commits the same
// shapes proof after proof, and a 9 GB request that finds its old block split by smaller allocations makes the pool
// map new physical memory (tens to hundreds of ms).  Reuse is ordered by ctx->stream like cudaFreeAsync / cudaMallocAsync
// on that stream would be.  The cache is dropped when an allocation fails (and on bsx_ctx_trim / destroy).
#define BSX_BIG_BLOCK_BYTES ((size_t)64 << 20)
void bsx_ctx_drop_cache(bsx_ctx *ctx) {
    for (auto &e : ctx->big_cache) cudaFreeAsync(e.first, ctx->stream);
    ctx->big_cache.clear();
    ctx->big_cache_bytes = 0;
}
int bsx_alloc_owned(bsx_ctx *ctx, u64 **p, size_t words) {
    *p = nullptr;
    if (words == 0) return BSX_OK;
    const size_t bytes = words * sizeof(u64);
    if (bytes >= BSX_BIG_BLOCK_BYTES) {
        for (size_t i = ctx->big_cache.size(); i-- > 0;)
            if (ctx->big_cache[i].second == bytes) {
                *p = ctx->big_cache[i].first;
                ctx->big_cache.erase(ctx->big_cache.begin() + (long)i);
                ctx->big_cache_bytes -= bytes;
                ctx->big_live[*p] = bytes;
                return BSX_OK;
            }
    }
    cudaError_t e = cudaMallocAsync((void **)p, bytes, ctx->stream);
    if (e == cudaErrorMemoryAllocation && !ctx->big_cache.empty()) {  // give the cached blocks back and try again
        cudaGetLastError();
        bsx_ctx_drop_cache(ctx);
        cudaStreamSynchronize(ctx->stream);
        e = cudaMallocAsync((void **)p, bytes, ctx->stream);
    }
    if (e != cudaSuccess) {
        cudaGetLastError();
        *p = nullptr;
        return bsx_fail(ctx, e == cudaErrorMemoryAllocation ? BSX_ERR_OOM : BSX_ERR_CUDA,
                        "device allocation of %zu bytes failed: %s", bytes, cudaGetErrorString(e));
    }
    if (bytes >= BSX_BIG_BLOCK_BYTES) ctx->big_live[*p] = bytes;
    return BSX_OK;
}
int bsx_alloc(bsx_ctx *ctx, u64 **p, size_t words) {
    TRY(bsx_alloc_owned(ctx, p, words));
    if (*p && ctx->scope_depth > 0) ctx->scratch.emplace_back(*p, ctx->scratch_seq++);
    return BSX_OK;
}
static void release_block(bsx_ctx *ctx, u64 *p) {
    auto it = ctx->big_live.find(p);
    if (it == ctx->big_live.end()) {
        cudaFreeAsync(p, ctx->stream);
        return;
    }
    ctx->big_cache.emplace_back(p, it->second);
    ctx->big_cache_bytes += it->second;
    ctx->big_live.erase(it);
}
void bsx_free(bsx_ctx *ctx, u64 *p) {
    if (!p) return;
    for (size_t i = ctx->scratch.size(); i-- > 0;)
        if (ctx->scratch[i].first == p) {
            ctx->scratch.erase(ctx->scratch.begin() + (long)i);
            break;
        }
    release_block(ctx, p);
}
// stream-ordered frees: everything the call enqueued before this point has the buffers it needs
BsxScope::~BsxScope() {
    if (!ctx) return;
    for (size_t i = ctx->scratch.size(); i-- > 0;)
        if (ctx->scratch[i].second >= first) {
            release_block(ctx, ctx->scratch[i].first);
            ctx->scratch.erase(ctx->scratch.begin() + (long)i);
        }
    ctx->scope_depth--;
}

int bsx_copy_in(bsx_ctx *ctx, u64 *dst_dev, const u64 *src, size_t words, int mem) {
    if (words == 0) return BSX_OK;
    CU(ctx, cudaMemcpyAsync(dst_dev, src, words * sizeof(u64),
                            mem == BSX_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, ctx->stream));
    return BSX_OK;
}
int bsx_copy_out(bsx_ctx *ctx, u64 *dst, const u64 *src_dev, size_t words, int mem) {
    if (words == 0) return BSX_OK;
    CU(ctx, cudaMemcpyAsync(dst, src_dev, words * sizeof(u64),
                            mem == BSX_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, ctx->stream));
    return BSX_OK;
}

// powers of omega_{2^log_n} (or its inverse), two-level: w^e = hi[e >> 12] * lo[e & 4095]
int bsx_pow_table(bsx_ctx *ctx, int log_n, int inverse, PowTable *out) {
    int key = log_n * 2 + (inverse ? 1 : 0);
    auto it = ctx->pow_tables.find(key);
    if (it != ctx->pow_tables.end()) {
        *out = it->second;
        return BSX_OK;
    }
    u64 w = gl_host_root(log_n);
    if (inverse) w = gl_host_inv(w);
    u64 *lo, *hi;
    TRY(bsx_alloc_owned(ctx, &lo, 4096));
    TRY(bsx_alloc_owned(ctx, &hi, 4096));
    launch_fill_powers(lo, w, 1, 4096, ctx->stream);
    launch_fill_powers(hi, gl_host_pow(w, 4096), 1, 4096, ctx->stream);
    CU(ctx, cudaGetLastError());
    PowTable t = {lo, hi};
    ctx->pow_tables[key] = t;
    *out = t;
    return BSX_OK;
}

void bsx_split(int log_n, int *a, int *b) {
    if (log_n <= ntt_max_single_pass_log()) {
        *a = log_n;
        *b = 0;
    } else {
        *b = log_n / 2;
        *a = log_n - *b;
    }
}

// coset pre-factor tables for g_k = shift * omega_{2^(log_n + r_log)}^k, k < 2^r_log
int bsx_coset_tables(bsx_ctx *ctx, int log_n, int r_log, u64 shift, const u64 **pre_hi, const u64 **pre_lo) {
    CosetKey key = {log_n, r_log, shift};
    auto it = ctx->coset_tables.find(key);
    if (it != ctx->coset_tables.end()) {
        *pre_hi = it->second.first;
        *pre_lo = it->second.second;
        return BSX_OK;
    }
    int a, b;
    bsx_split(log_n, &a, &b);
    int nc = 1 << r_log;
    u64 *hi, *lo;
    TRY(bsx_alloc_owned(ctx, &hi, (size_t)2 * nc << a));  // natural order, then the same rows bit-reversed
    TRY(bsx_alloc_owned(ctx, &lo, (size_t)nc << b));
    u64 w = gl_host_root(log_n + r_log);
    for (int k = 0; k < nc; k++) {
        u64 g = gl_host_mul(shift, gl_host_pow(w, (u64)k));
        launch_fill_powers(hi + ((size_t)k << a), gl_host_pow(g, (u64)1 << b), 1, (size_t)1 << a, ctx->stream);
        launch_fill_powers(lo + ((size_t)k << b), g, 1, (size_t)1 << b, ctx->stream);
    }
    // pass HI reads bit-reversed coefficient storage with consecutive lanes on consecutive words; the
    // bit-reversed copy makes its factor loads contiguous too (the natural table costs 32 sectors per warp)
    launch_bitrev_copy(hi, (size_t)1 << a, hi + ((size_t)nc << a), (size_t)1 << a, a, nc, ctx->stream);
    CU(ctx, cudaGetLastError());
    ctx->coset_tables[key] = std::make_pair((const u64 *)hi, (const u64 *)lo);
    *pre_hi = hi;
    *pre_lo = lo;
    return BSX_OK;
}

int bsx_post_table(bsx_ctx *ctx, NttPlan *plan, u64 shift, int n_cols) {
    plan->post_full = nullptr;
    int r_log = 0;
    while ((1 << r_log) < plan->n_cosets) r_log++;
    size_t words = (size_t)1 << (plan->log_n + r_log);
    // worth building only when many columns share it; bounded so that it never rivals the data
    if (plan->b == 0 || n_cols < 8 || words > ((size_t)1 << 28)) return BSX_OK;
    CosetKey key = {plan->log_n, r_log, plan->pre_hi ? shift : 0};
    if (plan->inverse) key.r_log |= 0x100;
    auto it = ctx->post_tables.find(key);
    if (it == ctx->post_tables.end()) {
        u64 *t;
        TRY(bsx_alloc_owned(ctx, &t, words));
        launch_ntt_post_table(t, plan->pre_hi ? plan->pre_lo : nullptr, plan->tw, plan->scale, plan->a, plan->b, r_log,
                              ctx->stream);
        CU(ctx, cudaGetLastError());
        it = ctx->post_tables.emplace(key, (const u64 *)t).first;
    }
    plan->post_full = it->second;
    return BSX_OK;
}

// ------------------------------------------------------------------------------------ context
unsigned long long g_bsx_launches = 0;
extern "C" uint64_t bsx_launch_count(void) { return g_bsx_launches; }

extern "C" const char *bsx_version(void) { return "bsx_b200 0.1 (sm_100a)"; }

extern "C" int32_t bsx_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

extern "C" int32_t bsx_ctx_create(int32_t device, bsx_ctx **out) {
    if (!out) return BSX_ERR_INVALID_ARG;
    *out = nullptr;
    int n = bsx_device_count();
    if (n <= 0 || device < 0 || device >= n) return BSX_ERR_CUDA;  // no device: fail loudly, no fallback
    if (cudaSetDevice(device) != cudaSuccess) return BSX_ERR_CUDA;
    bsx_ctx *ctx = new bsx_ctx();
    ctx->device = device;
    ctx->err[0] = 0;
    if (cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess) {
        delete ctx;
        return BSX_ERR_CUDA;
    }
    if (cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking) != cudaSuccess) {
        cudaStreamDestroy(ctx->stream);
        delete ctx;
        return BSX_ERR_CUDA;
    }
    cudaMemPool_t pool;
    if (cudaDeviceGetDefaultMemPool(&pool, device) == cudaSuccess) {
        unsigned long long thr = ~0ULL;  // keep freed blocks cached: repeated commits reuse them
        cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thr);
    }
    ntt_init_device();
    u64 *wf = nullptr, *wi = nullptr;
    if (bsx_alloc_owned(ctx, &wf, 4096) || bsx_alloc_owned(ctx, &wi, 4096)) {
        cudaStreamDestroy(ctx->stream);
        delete ctx;
        return BSX_ERR_OOM;
    }
    u64 w = gl_host_root(12);
    launch_fill_powers(wf, w, 1, 4096, ctx->stream);  // all 4096 powers: the radix stage groups index the full circle
    launch_fill_powers(wi, gl_host_inv(w), 1, 4096, ctx->stream);
    ctx->tables.w_fwd = wf;
    ctx->tables.w_inv = wi;
    if (cudaStreamSynchronize(ctx->stream) != cudaSuccess) {
        cudaStreamDestroy(ctx->stream);
        delete ctx;
        return BSX_ERR_CUDA;
    }
    *out = ctx;
    return BSX_OK;
}

extern "C" void bsx_ctx_destroy(bsx_ctx *ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    bsx_free(ctx, (u64 *)ctx->tables.w_fwd);
    bsx_free(ctx, (u64 *)ctx->tables.w_inv);
    for (auto &kv : ctx->pow_tables) {
        bsx_free(ctx, (u64 *)kv.second.lo);
        bsx_free(ctx, (u64 *)kv.second.hi);
    }
    for (auto &kv : ctx->post_tables) bsx_free(ctx, (u64 *)kv.second);
    for (auto &kv : ctx->coset_tables) {
        bsx_free(ctx, (u64 *)kv.second.first);
        bsx_free(ctx, (u64 *)kv.second.second);
    }
    bsx_ctx_drop_cache(ctx);
    cudaStreamSynchronize(ctx->stream);
    if (ctx->stage_ring) cudaFreeHost(ctx->stage_ring);
    for (auto &e : ctx->stage_done)
        if (e) cudaEventDestroy(e);
    cudaStreamDestroy(ctx->copy_stream);
    cudaStreamDestroy(ctx->stream);
    delete ctx;
}

extern "C" const char *bsx_last_error(const bsx_ctx *ctx) { return ctx ? ctx->err : "null context"; }
extern "C" void *bsx_ctx_stream(const bsx_ctx *ctx) { return ctx ? (void *)ctx->stream : nullptr; }
// bytes of the device's stream-ordered pool currently handed out (all contexts of this process on the device)
extern "C" int32_t bsx_ctx_mem_used(bsx_ctx *ctx, uint64_t *bytes) {
    if (!ctx || !bytes) return BSX_ERR_INVALID_ARG;
    CU(ctx, cudaSetDevice(ctx->device));
    CU(ctx, cudaStreamSynchronize(ctx->stream));
    cudaMemPool_t pool;
    CU(ctx, cudaDeviceGetDefaultMemPool(&pool, ctx->device));
    unsigned long long used = 0;
    CU(ctx, cudaMemPoolGetAttribute(pool, cudaMemPoolAttrUsedMemCurrent, &used));
    *bytes = used - ctx->big_cache_bytes;  // blocks parked for reuse are not in use
    return BSX_OK;
}
extern "C" int32_t bsx_ctx_trim(bsx_ctx *ctx) {
    if (!ctx) return BSX_ERR_INVALID_ARG;
    CU(ctx, cudaSetDevice(ctx->device));
    bsx_ctx_drop_cache(ctx);
    CU(ctx, cudaStreamSynchronize(ctx->stream));
    return BSX_OK;
}
extern "C" int32_t bsx_ctx_sync(bsx_ctx *ctx) {
    if (!ctx) return BSX_ERR_INVALID_ARG;
    CU(ctx, cudaStreamSynchronize(ctx->stream));
    return BSX_OK;
}
