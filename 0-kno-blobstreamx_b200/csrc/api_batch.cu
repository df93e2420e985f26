// api_batch.cu -- PolynomialBatch and MerkleTree entry points of the C ABI.
// Upstream: plonky2 fri/oracle.rs (from_values, from_coeffs, lde_values, get_lde_values),
// hash/merkle_tree.rs (MerkleTree::new, get, prove).
#include "api_internal.h"

#include <atomic>
#include <cstdlib>
#include <cstring>
#include <condition_variable>
#include <mutex>
#include <thread>
#include <sys/random.h>

static inline size_t n_digest_words(int log_leaves, int cap_height) {
    return 4 * 2 * (((size_t)1 << log_leaves) - ((size_t)1 << cap_height));
}

// ------------------------------------------------------------------------------------ MerkleTree
int bsx_tree_alloc(bsx_tree *t) {
    bsx_ctx *ctx = t->ctx;
    t->digests = nullptr;
    t->cap_dev = nullptr;
    TRY(bsx_alloc_owned(ctx, &t->cap_dev, (size_t)4 << t->cap_height));
    TRY(bsx_alloc_owned(ctx, &t->digests, n_digest_words(t->log_leaves, t->cap_height)));
    return BSX_OK;
}

// digests of the leaves [first, first + count) on stream st (the commit hashes finished LDE blocks on a second
// stream while the next block is still being transformed)
void bsx_tree_hash_leaves(bsx_tree *t, size_t first, size_t count, cudaStream_t st) {
    // cap_height == log2(leaves): the leaf digests are the cap (merkle_tree.rs fill_digests_buf)
    u64 *leaf_digests = t->cap_height == t->log_leaves ? t->cap_dev : t->digests;
    launch_leaf_hash(t->leaves + first, t->col_stride, t->leaf_len, count, leaf_digests + 4 * first, st);
}

int bsx_tree_finish(bsx_tree *t) {
    bsx_ctx *ctx = t->ctx;
    const size_t cap_words = (size_t)4 << t->cap_height;
    launch_merkle_tree(t->digests, t->cap_dev, t->log_leaves, t->cap_height, ctx->stream);
    CU(ctx, cudaGetLastError());
    t->cap_host.resize(cap_words);
    CU(ctx, cudaMemcpyAsync(t->cap_host.data(), t->cap_dev, cap_words * sizeof(u64), cudaMemcpyDeviceToHost,
                            ctx->stream));
    return BSX_OK;
}

int bsx_tree_build(bsx_tree *t, cudaEvent_t ev_after_leaf_hash) {
    TRY(bsx_tree_alloc(t));
    bsx_tree_hash_leaves(t, 0, (size_t)1 << t->log_leaves, t->ctx->stream);
    if (ev_after_leaf_hash) cudaEventRecord(ev_after_leaf_hash, t->ctx->stream);
    return bsx_tree_finish(t);
}

void bsx_tree_release(bsx_tree *t) {
    if (t->owns_leaves) bsx_free(t->ctx, (u64 *)t->leaves);
    bsx_free(t->ctx, t->digests);
    bsx_free(t->ctx, t->cap_dev);
    t->leaves = nullptr;
    t->digests = t->cap_dev = nullptr;
}

static int tree_check_shape(bsx_ctx *ctx, int log_leaves, int leaf_len, int cap_height) {
    if (log_leaves < 0 || log_leaves > 40 || leaf_len < 0)
        return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "bad tree shape: log_leaves=%d leaf_len=%d", log_leaves, leaf_len);
    if (cap_height < 0 || cap_height > log_leaves)  // upstream: assert in MerkleTree::new
        return bsx_fail(ctx, BSX_ERR_INVALID_ARG,
                        "cap_height=%d should be at most log2(leaves.len())=%d", cap_height, log_leaves);
    return BSX_OK;
}

extern "C" int32_t bsx_merkle_new(bsx_ctx *ctx, const uint64_t *leaves, int32_t mem, int32_t log_leaves,
                                  int32_t leaf_len, int32_t cap_height, bsx_tree **out) {
    if (!ctx || !out) return BSX_ERR_INVALID_ARG;
    BsxScope scope(ctx);
    *out = nullptr;
    TRY(tree_check_shape(ctx, log_leaves, leaf_len, cap_height));
    if (!leaves && leaf_len > 0) return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "leaves is NULL");
    CU(ctx, cudaSetDevice(ctx->device));
    const size_t n = (size_t)1 << log_leaves, words = n * (size_t)leaf_len;
    u64 *rowmajor = nullptr, *colmajor = nullptr;
    const u64 *src = (const u64 *)leaves;
    if (mem != BSX_DEVICE) {
        TRY(bsx_alloc(ctx, &rowmajor, words));
        TRY(bsx_copy_in(ctx, rowmajor, (const u64 *)leaves, words, mem));
        src = rowmajor;
    }
    TRY(bsx_alloc_owned(ctx, &colmajor, words));
    launch_transpose(src, colmajor, n, (size_t)leaf_len, ctx->stream);
    bsx_free(ctx, rowmajor);
    bsx_tree *t = new bsx_tree();
    t->ctx = ctx;
    t->log_leaves = log_leaves;
    t->leaf_len = leaf_len;
    t->cap_height = cap_height;
    t->leaves = colmajor;
    t->col_stride = n;
    t->owns_leaves = true;
    t->embedded = false;
    int rc = bsx_tree_build(t, nullptr);
    if (rc == BSX_OK && cudaStreamSynchronize(ctx->stream) != cudaSuccess)
        rc = bsx_fail(ctx, BSX_ERR_CUDA, "merkle build failed: %s", cudaGetErrorString(cudaGetLastError()));
    if (rc != BSX_OK) {
        bsx_tree_release(t);
        delete t;
        return rc;
    }
    *out = t;
    return BSX_OK;
}

extern "C" int32_t bsx_merkle_new_colmajor_device(bsx_ctx *ctx, const uint64_t *leaves_dev, size_t col_stride,
                                                  int32_t log_leaves, int32_t leaf_len, int32_t cap_height,
                                                  bsx_tree **out) {
    if (!ctx || !out) return BSX_ERR_INVALID_ARG;
    BsxScope scope(ctx);
    *out = nullptr;
    TRY(tree_check_shape(ctx, log_leaves, leaf_len, cap_height));
    if (col_stride < ((size_t)1 << log_leaves)) return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "col_stride too small");
    CU(ctx, cudaSetDevice(ctx->device));
    bsx_tree *t = new bsx_tree();
    t->ctx = ctx;
    t->log_leaves = log_leaves;
    t->leaf_len = leaf_len;
    t->cap_height = cap_height;
    t->leaves = (const u64 *)leaves_dev;
    t->col_stride = col_stride;
    t->owns_leaves = false;
    t->embedded = false;
    int rc = bsx_tree_build(t, nullptr);
    if (rc == BSX_OK && cudaStreamSynchronize(ctx->stream) != cudaSuccess)
        rc = bsx_fail(ctx, BSX_ERR_CUDA, "merkle build failed: %s", cudaGetErrorString(cudaGetLastError()));
    if (rc != BSX_OK) {
        bsx_tree_release(t);
        delete t;
        return rc;
    }
    *out = t;
    return BSX_OK;
}

extern "C" void bsx_merkle_free(bsx_tree *t) {
    if (!t || t->embedded) return;
    cudaSetDevice(t->ctx->device);
    bsx_tree_release(t);
    delete t;
}

extern "C" int32_t bsx_merkle_cap(const bsx_tree *t, uint64_t *out_host) {
    if (!t || !out_host) return BSX_ERR_INVALID_ARG;
    memcpy(out_host, t->cap_host.data(), t->cap_host.size() * sizeof(u64));
    return BSX_OK;
}

extern "C" int32_t bsx_merkle_shape(const bsx_tree *t, int32_t *log_leaves, int32_t *leaf_len, int32_t *cap_height) {
    if (!t) return BSX_ERR_INVALID_ARG;
    if (log_leaves) *log_leaves = t->log_leaves;
    if (leaf_len) *leaf_len = t->leaf_len;
    if (cap_height) *cap_height = t->cap_height;
    return BSX_OK;
}

extern "C" int32_t bsx_merkle_get(bsx_tree *t, const uint64_t *idx_host, int32_t k, uint64_t *out_host) {
    if (!t || k < 0 || (k > 0 && (!idx_host || !out_host))) return BSX_ERR_INVALID_ARG;
    bsx_ctx *ctx = t->ctx;
    BsxScope scope(ctx);
    if (k == 0 || t->leaf_len == 0) return BSX_OK;
    for (int i = 0; i < k; i++)
        if (idx_host[i] >> t->log_leaves)
            return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "leaf index %llu out of range", (unsigned long long)idx_host[i]);
    CU(ctx, cudaSetDevice(ctx->device));
    u64 *idx_dev, *rows_dev;
    TRY(bsx_alloc(ctx, &idx_dev, (size_t)k));
    TRY(bsx_alloc(ctx, &rows_dev, (size_t)k * t->leaf_len));
    CU(ctx, cudaMemcpyAsync(idx_dev, idx_host, (size_t)k * sizeof(u64), cudaMemcpyHostToDevice, ctx->stream));
    launch_gather_rows(t->leaves, t->col_stride, t->leaf_len, idx_dev, k, rows_dev, ctx->stream);
    CU(ctx, cudaMemcpyAsync(out_host, rows_dev, (size_t)k * t->leaf_len * sizeof(u64), cudaMemcpyDeviceToHost,
                            ctx->stream));
    bsx_free(ctx, idx_dev);
    bsx_free(ctx, rows_dev);
    CU(ctx, cudaStreamSynchronize(ctx->stream));
    return BSX_OK;
}

extern "C" int32_t bsx_merkle_prove(bsx_tree *t, uint64_t leaf_index, uint64_t *siblings_host) {
    if (!t) return BSX_ERR_INVALID_ARG;
    bsx_ctx *ctx = t->ctx;
    BsxScope scope(ctx);
    if (leaf_index >> t->log_leaves)
        return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "leaf index %llu out of range", (unsigned long long)leaf_index);
    int levels = t->log_leaves - t->cap_height;
    if (levels == 0) return BSX_OK;
    if (!siblings_host) return BSX_ERR_INVALID_ARG;
    CU(ctx, cudaSetDevice(ctx->device));
    u64 *sib;
    TRY(bsx_alloc(ctx, &sib, (size_t)4 * levels));
    launch_gather_siblings(t->digests, t->log_leaves, levels, leaf_index, sib, ctx->stream);
    CU(ctx, cudaMemcpyAsync(siblings_host, sib, (size_t)4 * levels * sizeof(u64), cudaMemcpyDeviceToHost, ctx->stream));
    bsx_free(ctx, sib);
    CU(ctx, cudaStreamSynchronize(ctx->stream));
    return BSX_OK;
}

// level-order -> plonky2's recursive `digests` layout (merkle_tree.rs fill_subtree):
// left recursive buffer || left child digest || right child digest || right recursive buffer.
static void to_plonky2_layout(const u64 *lv, size_t n, int height, size_t node, u64 *buf, size_t buf_len) {
    if (buf_len == 0) return;
    size_t half = buf_len / 2, off = 0;
    for (int k = 0; k < height - 1; k++) off += n >> k;
    memcpy(buf + 4 * (half - 1), lv + 4 * (off + 2 * node), 4 * sizeof(u64));
    memcpy(buf + 4 * half, lv + 4 * (off + 2 * node + 1), 4 * sizeof(u64));
    to_plonky2_layout(lv, n, height - 1, 2 * node, buf, half - 1);
    to_plonky2_layout(lv, n, height - 1, 2 * node + 1, buf + 4 * (half + 1), half - 1);
}

extern "C" int32_t bsx_merkle_digests(bsx_tree *t, uint64_t *out_host) {
    if (!t) return BSX_ERR_INVALID_ARG;
    bsx_ctx *ctx = t->ctx;
    BsxScope scope(ctx);
    size_t words = n_digest_words(t->log_leaves, t->cap_height);
    if (words == 0) return BSX_OK;
    if (!out_host) return BSX_ERR_INVALID_ARG;
    CU(ctx, cudaSetDevice(ctx->device));
    std::vector<u64> lv(words);
    CU(ctx, cudaMemcpyAsync(lv.data(), t->digests, words * sizeof(u64), cudaMemcpyDeviceToHost, ctx->stream));
    CU(ctx, cudaStreamSynchronize(ctx->stream));
    int height = t->log_leaves - t->cap_height;
    size_t sub = ((size_t)2 << height) - 2, n = (size_t)1 << t->log_leaves;
    for (size_t c = 0; c < ((size_t)1 << t->cap_height); c++)
        to_plonky2_layout(lv.data(), n, height, c, (u64 *)out_host + 4 * c * sub, sub);
    return BSX_OK;
}

// ------------------------------------------------------------------------------------ PolynomialBatch
struct PhaseTimer {
    cudaEvent_t ev[6];
    bool on;
    PhaseTimer(bool enable) : on(enable) {
        if (on)
            for (auto &e : ev) cudaEventCreate(&e);
    }
    ~PhaseTimer() {
        if (on)
            for (auto &e : ev) cudaEventDestroy(e);
    }
    void mark(int i, cudaStream_t st) {
        if (on) cudaEventRecord(ev[i], st);
    }
    float ms(int i, int j) {
        float t = 0;
        cudaEventElapsedTime(&t, ev[i], ev[j]);
        return t;
    }
};

// iNTT (unless the input already holds coefficients) and the coset LDE of n_cols columns: the common front
// half of PolynomialBatch::from_values / from_coeffs, also used per rank by the row-sharded commit
// (api_shard.cu), where `dests` redirects the final LDE words to the owning ranks' row slabs.
// coeffs: n_cols x 2^log_n in place (left bit-reversed after the inverse transform);
// lde: n_cols columns of stride lde_stride (the result, or only the inter-pass scratch when dests is set).
int bsx_commit_ifft(bsx_ctx *ctx, u64 *coeffs, int log_n, int n_cols) {
    const size_t n = (size_t)1 << log_n;
    NttPlan plan;
    plan.n_dest = 0;
    plan.log_n = log_n;
    bsx_split(log_n, &plan.a, &plan.b);
    // PolynomialValues::ifft per column, in place; result is stored bit-reversed
    plan.inverse = 1;
    plan.src_bitrev = 0;
    plan.n_cosets = 1;
    plan.pre_hi = plan.pre_lo = nullptr;
    TRY(bsx_pow_table(ctx, log_n, 1, &plan.tw));
    plan.scale = gl_host_inv((u64)n % GL_P);
    TRY(bsx_post_table(ctx, &plan, 0, n_cols));
    launch_ntt(ctx->tables, plan, coeffs, n, coeffs, n, n_cols, ctx->stream);
    CU(ctx, cudaGetLastError());
    return BSX_OK;
}

// lde_values: zero-pad + coset FFT(shift = 7) == 2^rate_bits coset transforms of size n
int bsx_commit_lde(bsx_ctx *ctx, const u64 *coeffs, bool coeffs_bitrev, int log_n, int n_cols, int rate_bits, u64 *lde,
                   size_t lde_stride, const BsxDests *dests) {
    const size_t n = (size_t)1 << log_n;
    NttPlan plan;
    plan.n_dest = 0;
    plan.log_n = log_n;
    bsx_split(log_n, &plan.a, &plan.b);
    plan.inverse = 0;
    plan.src_bitrev = coeffs_bitrev ? 1 : 0;
    plan.n_cosets = 1 << rate_bits;
    TRY(bsx_coset_tables(ctx, log_n, rate_bits, 7, &plan.pre_hi, &plan.pre_lo));
    TRY(bsx_pow_table(ctx, log_n, 0, &plan.tw));
    plan.scale = 1;
    TRY(bsx_post_table(ctx, &plan, 7, n_cols));
    if (dests) {
        plan.n_dest = dests->n;
        plan.dest_lc = dests->lc;
        for (int i = 0; i < dests->n; i++) plan.dest[i] = dests->ptr[i];
    }
    launch_ntt(ctx->tables, plan, coeffs, n, lde, lde_stride, n_cols, ctx->stream);
    CU(ctx, cudaGetLastError());
    return BSX_OK;
}

int bsx_commit_transforms(bsx_ctx *ctx, u64 *coeffs, int log_n, int n_cols, int rate_bits, bool input_is_coeffs,
                          u64 *lde, size_t lde_stride, const BsxDests *dests, cudaEvent_t ev_after_ifft) {
    if (!input_is_coeffs) TRY(bsx_commit_ifft(ctx, coeffs, log_n, n_cols));
    if (ev_after_ifft) cudaEventRecord(ev_after_ifft, ctx->stream);
    return bsx_commit_lde(ctx, coeffs, !input_is_coeffs, log_n, n_cols, rate_bits, lde, lde_stride, dests);
}

// ---- host witness sources -------------------------------------------------------------------------------------
// A witness arrives either as one column-major buffer or as one pointer per column (`Vec<PolynomialValues<F>>`:
// every column its own heap allocation).  Pinned contiguous memory is copied by the DMA engine directly.  Pageable
// memory (what a Rust Vec is) would make cudaMemcpyAsync stage through the driver at ~10 GB/s on the calling
// thread; instead a few host threads copy slot-sized pieces into a pinned ring owned by the context while the DMA
// engine drains the previous slots, and the transforms of the finished column chunks run underneath.
struct HostColumns {
    const u64 *contig;         // n_cols x n, or null
    const u64 *const *cols;    // n_cols pointers, or null
    size_t n;
    const u64 *col(int c) const { return cols ? cols[c] : contig + (size_t)c * n; }
};

static bool host_is_pinned(const void *p) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return a.type == cudaMemoryTypeHost;
}

// Host threads that copy pageable columns into the pinned ring.  Default: the host's hardware threads shared between
// the ranks of this host (LOCAL_WORLD_SIZE, set by torchrun / most launchers; 1 otherwise), at most 8 per rank: eight
// ranks with eight threads each on a 16-core box spend their time in the scheduler, not in memcpy.  BSX_COPY_THREADS
// overrides.
static int stage_threads() {
    static int t = 0;
    if (t == 0) {
        const char *e = getenv("BSX_COPY_THREADS");
        int hw = (int)std::thread::hardware_concurrency();
        const char *lw = getenv("LOCAL_WORLD_SIZE");
        int ranks = lw ? atoi(lw) : 1;
        if (ranks < 1) ranks = 1;
        t = e ? atoi(e) : (hw >= 16 ? (hw / ranks < 8 ? hw / ranks : 8) : hw >= 4 ? hw / 2 / ranks : 1);
        if (t < 1) t = 1;
        if (t > 32) t = 32;
    }
    return t;
}

// Streaming copy: the staged words are read once more, by the DMA engine, never by this core, so they should not
// displace the caller's working set nor cost a read-for-ownership of the destination lines.  glibc's memcpy only
// switches to non-temporal stores above a per-thread threshold that a 1 MiB piece stays below.
#if defined(__x86_64__)
#include <immintrin.h>
__attribute__((target("avx2"))) static void copy_stream_avx2(u64 *dst, const u64 *src, size_t words) {
    size_t i = 0;
    while (i < words && ((uintptr_t)(dst + i) & 31)) {
        dst[i] = src[i];
        i++;
    }
    for (; i + 16 <= words; i += 16) {
        __m256i a = _mm256_loadu_si256((const __m256i *)(src + i)), b = _mm256_loadu_si256((const __m256i *)(src + i + 4));
        __m256i c = _mm256_loadu_si256((const __m256i *)(src + i + 8)), d = _mm256_loadu_si256((const __m256i *)(src + i + 12));
        _mm256_stream_si256((__m256i *)(dst + i), a);
        _mm256_stream_si256((__m256i *)(dst + i + 4), b);
        _mm256_stream_si256((__m256i *)(dst + i + 8), c);
        _mm256_stream_si256((__m256i *)(dst + i + 12), d);
    }
    for (; i < words; i++) dst[i] = src[i];
    _mm_sfence();
}
#endif
static void copy_words(u64 *dst, const u64 *src, size_t words) {
#if defined(__x86_64__)
    static const int nt = (getenv("BSX_COPY_NT") ? atoi(getenv("BSX_COPY_NT")) : 1) && __builtin_cpu_supports("avx2");
    if (nt) {
        copy_stream_avx2(dst, src, words);
        return;
    }
#endif
    memcpy(dst, src, words * sizeof(u64));
}

// A small persistent pool per process: the ring is filled slot by slot (~35 slots per 1.13 GB witness), and creating
// the threads anew for every slot cost more than some of the copies.
namespace {
struct CopyJob {
    u64 *dst;
    const HostColumns *src;
    int c0, cc;
    size_t per_col, total, piece;
    std::atomic<size_t> next{0};
    std::atomic<int> left{0};
};
struct CopyPool {
    std::mutex m;
    std::condition_variable cv, done_cv;
    std::vector<std::thread> workers;
    CopyJob *job = nullptr;
    unsigned long long generation = 0;
    bool stop = false;
    static void run(CopyJob *j) {
        for (;;) {
            size_t t = j->next.fetch_add(1);
            if (t >= j->total) return;
            size_t c = t / j->per_col, off = (t % j->per_col) * j->piece;
            size_t len = j->src->n - off < j->piece ? j->src->n - off : j->piece;
            copy_words(j->dst + c * j->src->n + off, j->src->col(j->c0 + (int)c) + off, len);
        }
    }
    void worker() {
        unsigned long long seen = 0;
        for (;;) {
            CopyJob *j;
            {
                std::unique_lock<std::mutex> lk(m);
                cv.wait(lk, [&] { return stop || generation != seen; });
                if (stop) return;
                seen = generation;
                j = job;
            }
            run(j);
            if (j->left.fetch_sub(1) == 1) {
                std::lock_guard<std::mutex> lk(m);
                done_cv.notify_all();
            }
        }
    }
    void ensure(int n) {
        try {
            while ((int)workers.size() < n) workers.emplace_back([this] { worker(); });
        } catch (...) {
            // thread creation refused (resource limits): the workers that exist and the caller finish the copy
        }
    }
    void copy(CopyJob &j, int threads) {
        ensure(threads - 1);
        j.left.store((int)workers.size() + 1);  // every worker sees the new generation and takes part
        {
            std::lock_guard<std::mutex> lk(m);
            job = &j;
            generation++;
        }
        cv.notify_all();
        run(&j);
        if (j.left.fetch_sub(1) != 1) {
            std::unique_lock<std::mutex> lk(m);
            done_cv.wait(lk, [&] { return j.left.load() == 0; });
        }
    }
    ~CopyPool() {
        {
            std::lock_guard<std::mutex> lk(m);
            stop = true;
        }
        cv.notify_all();
        for (auto &t : workers) t.join();
    }
};
CopyPool &copy_pool() {
    static CopyPool p;
    return p;
}
std::mutex g_copy_pool_user;  // one staged copy at a time per process: contexts on other threads wait their turn
}  // namespace

// columns [c0, c0 + cc) -> dst (pinned), split into 1 MiB pieces taken by `threads` workers
static void stage_copy(u64 *dst, const HostColumns &src, int c0, int cc, int threads) {
    CopyJob j;
    j.dst = dst;
    j.src = &src;
    j.c0 = c0;
    j.cc = cc;
    j.piece = (size_t)1 << 17;  // words
    j.per_col = (src.n + j.piece - 1) / j.piece;
    j.total = j.per_col * (size_t)cc;
    if ((size_t)threads > j.total) threads = (int)j.total;
    if (threads <= 1) {
        CopyPool::run(&j);
        return;
    }
    std::lock_guard<std::mutex> user(g_copy_pool_user);
    copy_pool().copy(j, threads);
}

static int stage_ring_reserve(bsx_ctx *ctx, size_t slot_words) {
    if (ctx->stage_slot_words >= slot_words) return BSX_OK;
    if (ctx->stage_ring) {
        cudaStreamSynchronize(ctx->copy_stream);
        cudaFreeHost(ctx->stage_ring);
        ctx->stage_ring = nullptr;
        ctx->stage_slot_words = 0;
    }
    cudaError_t e = cudaHostAlloc((void **)&ctx->stage_ring, slot_words * BSX_STAGE_SLOTS * sizeof(u64), cudaHostAllocDefault);
    if (e != cudaSuccess) {
        cudaGetLastError();
        ctx->stage_ring = nullptr;
        return bsx_fail(ctx, BSX_ERR_OOM, "pinned staging ring of %zu bytes: %s", slot_words * BSX_STAGE_SLOTS * sizeof(u64),
                        cudaGetErrorString(e));
    }
    ctx->stage_slot_words = slot_words;
    for (auto &ev : ctx->stage_done)
        if (!ev) cudaEventCreateWithFlags(&ev, cudaEventDisableTiming);
    return BSX_OK;
}

// Enqueues on copy_stream the upload of columns [c0, c0 + cc) to dst (device, column stride n).  `seq` counts the
// ring slots used so far in this call.
static int upload_columns(bsx_ctx *ctx, const HostColumns &src, bool direct, int c0, int cc, u64 *dst, size_t *seq) {
    const size_t n = src.n;
    if (direct) {
        if (src.contig) {
            CU(ctx, cudaMemcpyAsync(dst, src.contig + (size_t)c0 * n, (size_t)cc * n * sizeof(u64), cudaMemcpyHostToDevice,
                                    ctx->copy_stream));
        } else {
            for (int c = 0; c < cc; c++)
                CU(ctx, cudaMemcpyAsync(dst + (size_t)c * n, src.cols[c0 + c], n * sizeof(u64), cudaMemcpyHostToDevice,
                                        ctx->copy_stream));
        }
        return BSX_OK;
    }
    const int per_slot = (int)(ctx->stage_slot_words / n);
    for (int c = 0; c < cc; c += per_slot) {
        const int k = cc - c < per_slot ? cc - c : per_slot;
        const int slot = (int)(*seq % BSX_STAGE_SLOTS);
        u64 *buf = ctx->stage_ring + (size_t)slot * ctx->stage_slot_words;
        if (*seq >= BSX_STAGE_SLOTS) CU(ctx, cudaEventSynchronize(ctx->stage_done[slot]));  // its previous DMA is over
        stage_copy(buf, src, c0 + c, k, stage_threads());
        CU(ctx, cudaMemcpyAsync(dst + (size_t)c * n, buf, (size_t)k * n * sizeof(u64), cudaMemcpyHostToDevice,
                                ctx->copy_stream));
        CU(ctx, cudaEventRecord(ctx->stage_done[slot], ctx->copy_stream));
        (*seq)++;
    }
    return BSX_OK;
}

static int batch_build(bsx_ctx *ctx, const u64 *input, const u64 *const *col_ptrs, int mem, int log_n, int n_cols, int rate_bits, int cap_height,
                       int blinding, const u64 *salt, bool input_is_coeffs, bsx_batch **out, bsx_timing *timing) {
    if (!ctx || !out) return BSX_ERR_INVALID_ARG;
    BsxScope scope(ctx);
    *out = nullptr;
    if (log_n < 0 || log_n > 24)
        return bsx_fail(ctx, BSX_ERR_UNSUPPORTED, "degree_log=%d outside the supported range [0, 24]", log_n);
    if (rate_bits < 0 || rate_bits > 8 || log_n + rate_bits > 32)
        return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "bad rate_bits=%d", rate_bits);
    if (n_cols < 0 || n_cols > 65535) return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "bad n_cols=%d", n_cols);
    const int log_nb = log_n + rate_bits;
    if (cap_height < 0 || cap_height > log_nb)
        return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "cap_height=%d should be at most log2(leaves.len())=%d",
                        cap_height, log_nb);
    if (!input && !col_ptrs && n_cols > 0) return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "input is NULL");
    if (col_ptrs) {
        if (mem != BSX_HOST) return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "column pointers must be host memory");
        for (int c = 0; c < n_cols; c++)
            if (!col_ptrs[c] && log_n >= 0) return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "column %d is NULL", c);
    }
    CU(ctx, cudaSetDevice(ctx->device));
    const size_t n = (size_t)1 << log_n, nb = (size_t)1 << log_nb;
    const int n_salt = blinding ? BSX_SALT_SIZE : 0;
    PhaseTimer pt(timing != nullptr);

    bsx_batch *bt = new bsx_batch();
    bt->ctx = ctx;
    bt->log_n = log_n;
    bt->n_cols = n_cols;
    bt->n_salt = n_salt;
    bt->rate_bits = rate_bits;
    bt->coeffs = bt->lde = nullptr;
    bt->tree.ctx = ctx;
    bt->tree.leaves = nullptr;
    bt->tree.digests = bt->tree.cap_dev = nullptr;
    bt->tree.owns_leaves = false;
    bt->tree.embedded = true;
    int rc = BSX_OK;
    bool pipelined = false;
    float piped_ifft_ms = 0, piped_lde_ms = 0;
    do {
        pt.mark(0, ctx->stream);
        if ((rc = bsx_alloc_owned(ctx, &bt->coeffs, n * n_cols))) break;
        if ((rc = bsx_alloc_owned(ctx, &bt->lde, nb * (size_t)(n_cols + n_salt)))) break;
        // salt columns first: nothing below depends on them except the leaf hashing
        if (n_salt) {
            u64 *dst = bt->lde + nb * (size_t)n_cols;
            if (salt) {
                u64 *tmp = nullptr;
                const u64 *src = salt;
                if (mem != BSX_DEVICE) {
                    if ((rc = bsx_alloc(ctx, &tmp, nb * n_salt))) break;
                    if ((rc = bsx_copy_in(ctx, tmp, salt, nb * n_salt, mem))) break;
                    src = tmp;
                }
                launch_bitrev_copy(src, nb, dst, nb, log_nb, n_salt, ctx->stream);
                bsx_free(ctx, tmp);
            } else {
                // F::rand_vec upstream (a CSPRNG): a fresh 256-bit ChaCha20 key from the kernel's entropy pool per batch
                u32 key[8];
                size_t got = 0;
                while (got < sizeof(key)) {
                    ssize_t r = getrandom((char *)key + got, sizeof(key) - got, 0);
                    if (r <= 0) break;
                    got += (size_t)r;
                }
                if (got != sizeof(key)) {
                    rc = bsx_fail(ctx, BSX_ERR_UNSUPPORTED, "getrandom failed: no entropy for the blinding salts");
                    break;
                }
                launch_fill_random(dst, nb * n_salt, key, 0, ctx->stream);
                memset(key, 0, sizeof(key));
            }
        }
        const HostColumns host = {col_ptrs ? nullptr : input, col_ptrs, n};
        if (mem == BSX_HOST && n_cols >= 16 && log_n >= 13) {
            // Host witness: the copy is pipelined with the transforms in column chunks (copy engine and SMs
            // overlap; at the headline 20 ms of PCIe time hide behind 36 ms of NTT).  Chunks keep >= 8 columns
            // so that every launch still shares the full post-factor table.
            const int chunk = n_cols / 8 >= 8 ? (n_cols + 7) / 8 : 8;
            // Only the first chunk's copy is exposed, so the schedule starts small (8, then 12 columns) and
            // reaches the full chunk once the transforms of a chunk outlast the copy of the next one.
            std::vector<int> cstart;  // chunk k covers columns [cstart[k], cstart[k + 1])
            {
                int c = 0;
                // BSX_CHUNK_SCHEDULE="4,8,12,..." (measurement aid): explicit chunk sizes; what they leave goes in default chunks
                if (const char *env = getenv("BSX_CHUNK_SCHEDULE")) {
                    for (const char *q = env; *q && c < n_cols;) {
                        int r = atoi(q);
                        if (r < 1) break;
                        if (r > n_cols - c) r = n_cols - c;
                        cstart.push_back(c);
                        c += r;
                        while (*q && *q != ',') q++;
                        if (*q == ',') q++;
                    }
                } else if (n_cols >= 64 && chunk > 12)
                    for (int r : {8, 12}) {
                        cstart.push_back(c);
                        c += r;
                    }
                while (c < n_cols) {
                    cstart.push_back(c);
                    c += chunk;
                }
                if (!getenv("BSX_CHUNK_SCHEDULE") && cstart.size() > 1 && n_cols - cstart.back() < 8) cstart.pop_back();  // no tile-starved tail
                cstart.push_back(n_cols);
            }
            const int n_chunks = (int)cstart.size() - 1;
            // pinned contiguous memory goes straight to the DMA engine; anything else through the staging ring
            const bool direct = !col_ptrs && host_is_pinned(input);
            if (!direct) {
                size_t per_slot = ((size_t)4 << 20) / n;  // ~32 MiB slots, at least one column
                if (per_slot < 1) per_slot = 1;
                if (per_slot > (size_t)chunk) per_slot = (size_t)chunk;
                if ((rc = stage_ring_reserve(ctx, per_slot * n))) break;
            }
            std::vector<cudaEvent_t> ev((size_t)n_chunks * 4 + 1);
            for (auto &e : ev) cudaEventCreateWithFlags(&e, pt.on ? cudaEventDefault : cudaEventDisableTiming);
            cudaEventRecord(ev[(size_t)n_chunks * 4], ctx->stream);  // the buffers exist (stream-ordered allocation)
            cudaStreamWaitEvent(ctx->copy_stream, ev[(size_t)n_chunks * 4], 0);
            size_t seq = 0;
            auto upload = [&](int k) -> int {
                const int c0 = cstart[k], cc = cstart[k + 1] - c0;
                int r = upload_columns(ctx, host, direct, c0, cc, bt->coeffs + (size_t)c0 * n, &seq);
                if (r == BSX_OK) cudaEventRecord(ev[4 * k], ctx->copy_stream);
                return r;
            };
            // direct: all copies are queued up front.  staged: chunk k + 1 is copied by the host threads after the
            // transforms of chunk k have been queued, so the GPU always has work while the host copies
            if (direct)
                for (int k = 0; k < n_chunks && rc == BSX_OK; k++) rc = upload(k);
            else
                rc = upload(0);
            for (int k = 0; k < n_chunks && rc == BSX_OK; k++) {
                const int c0 = cstart[k], cc = cstart[k + 1] - c0;
                cudaStreamWaitEvent(ctx->stream, ev[4 * k], 0);
                cudaEventRecord(ev[4 * k + 1], ctx->stream);
                if (k == 0) pt.mark(1, ctx->stream);
                rc = bsx_commit_transforms(ctx, bt->coeffs + (size_t)c0 * n, log_n, cc, rate_bits, input_is_coeffs,
                                           bt->lde + (size_t)c0 * nb, nb, nullptr, ev[4 * k + 2]);
                cudaEventRecord(ev[4 * k + 3], ctx->stream);
                if (!direct && rc == BSX_OK && k + 1 < n_chunks) rc = upload(k + 1);
            }
            pt.mark(2, ctx->stream);
            cudaError_t e = cudaStreamSynchronize(ctx->stream);  // `input` may be released by the caller after return
            if (rc == BSX_OK && e != cudaSuccess) rc = bsx_fail(ctx, BSX_ERR_CUDA, "pipelined copy failed: %s", cudaGetErrorString(e));
            if (pt.on && rc == BSX_OK) {
                for (int k = 0; k < n_chunks; k++) {
                    float a = 0, b = 0;
                    cudaEventElapsedTime(&a, ev[4 * k + 1], ev[4 * k + 2]);
                    cudaEventElapsedTime(&b, ev[4 * k + 2], ev[4 * k + 3]);
                    piped_ifft_ms += a;
                    piped_lde_ms += b;
                }
            }
            for (auto &x : ev) cudaEventDestroy(x);
            if (rc) break;
            pipelined = true;
        } else {
            if (col_ptrs) {
                for (int c = 0; c < n_cols && rc == BSX_OK; c++) rc = bsx_copy_in(ctx, bt->coeffs + (size_t)c * n, col_ptrs[c], n, mem);
                if (rc) break;
            } else if ((rc = bsx_copy_in(ctx, bt->coeffs, input, n * n_cols, mem)))
                break;
            pt.mark(1, ctx->stream);
            if ((rc = bsx_commit_transforms(ctx, bt->coeffs, log_n, n_cols, rate_bits, input_is_coeffs, bt->lde, nb,
                                            nullptr, pt.on ? pt.ev[2] : nullptr)))
                break;
        }
        bt->coeffs_bitrev = !input_is_coeffs;
        if (cudaGetLastError() != cudaSuccess) {
            rc = bsx_fail(ctx, BSX_ERR_CUDA, "NTT launch failed");
            break;
        }
        pt.mark(3, ctx->stream);
        bt->tree.log_leaves = log_nb;
        bt->tree.leaf_len = n_cols + n_salt;
        bt->tree.cap_height = cap_height;
        bt->tree.leaves = bt->lde;
        bt->tree.col_stride = nb;
        if ((rc = bsx_tree_build(&bt->tree, pt.on ? pt.ev[4] : nullptr))) break;
        pt.mark(5, ctx->stream);
        cudaError_t e = cudaStreamSynchronize(ctx->stream);
        if (e != cudaSuccess) {
            rc = bsx_fail(ctx, BSX_ERR_CUDA, "commit failed: %s", cudaGetErrorString(e));
            break;
        }
    } while (0);
    if (rc != BSX_OK) {
        bsx_tree_release(&bt->tree);
        bsx_free(ctx, bt->coeffs);
        bsx_free(ctx, bt->lde);
        delete bt;
        return rc;
    }
    if (timing) {
        timing->h2d_ms = pt.ms(0, 1);  // pipelined host path: only the exposed wait for the first chunk
        timing->ifft_ms = pipelined ? piped_ifft_ms : pt.ms(1, 2);
        timing->lde_ms = pipelined ? piped_lde_ms : pt.ms(2, 3);
        timing->leaf_hash_ms = pt.ms(3, 4);
        timing->tree_ms = pt.ms(4, 5);
        timing->total_ms = pt.ms(0, 5);
    }
    *out = bt;
    return BSX_OK;
}

extern "C" int32_t bsx_batch_from_values(bsx_ctx *ctx, const uint64_t *values, int32_t mem, int32_t log_n,
                                         int32_t n_cols, int32_t rate_bits, int32_t cap_height, int32_t blinding,
                                         const uint64_t *salt, bsx_batch **out, bsx_timing *timing) {
    return batch_build(ctx, (const u64 *)values, nullptr, mem, log_n, n_cols, rate_bits, cap_height, blinding,
                       (const u64 *)salt, false, out, timing);
}
extern "C" int32_t bsx_batch_from_values_cols(bsx_ctx *ctx, const uint64_t *const *cols_host, int32_t log_n, int32_t n_cols,
                                              int32_t rate_bits, int32_t cap_height, int32_t blinding,
                                              const uint64_t *salt_host, bsx_batch **out, bsx_timing *timing) {
    if (!cols_host && n_cols > 0) return BSX_ERR_INVALID_ARG;
    return batch_build(ctx, nullptr, (const u64 *const *)cols_host, BSX_HOST, log_n, n_cols, rate_bits, cap_height,
                       blinding, (const u64 *)salt_host, false, out, timing);
}
extern "C" int32_t bsx_batch_from_coeffs_cols(bsx_ctx *ctx, const uint64_t *const *cols_host, int32_t log_n, int32_t n_cols,
                                              int32_t rate_bits, int32_t cap_height, int32_t blinding,
                                              const uint64_t *salt_host, bsx_batch **out, bsx_timing *timing) {
    if (!cols_host && n_cols > 0) return BSX_ERR_INVALID_ARG;
    return batch_build(ctx, nullptr, (const u64 *const *)cols_host, BSX_HOST, log_n, n_cols, rate_bits, cap_height,
                       blinding, (const u64 *)salt_host, true, out, timing);
}
extern "C" int32_t bsx_batch_from_coeffs(bsx_ctx *ctx, const uint64_t *coeffs, int32_t mem, int32_t log_n,
                                         int32_t n_cols, int32_t rate_bits, int32_t cap_height, int32_t blinding,
                                         const uint64_t *salt, bsx_batch **out, bsx_timing *timing) {
    return batch_build(ctx, (const u64 *)coeffs, nullptr, mem, log_n, n_cols, rate_bits, cap_height, blinding,
                       (const u64 *)salt, true, out, timing);
}

extern "C" void bsx_batch_free(bsx_batch *b) {
    if (!b) return;
    cudaSetDevice(b->ctx->device);
    bsx_tree_release(&b->tree);
    bsx_free(b->ctx, b->coeffs);
    bsx_free(b->ctx, b->lde);
    delete b;
}

extern "C" int32_t bsx_batch_cap(const bsx_batch *b, uint64_t *out_host) {
    return b ? bsx_merkle_cap(&b->tree, out_host) : BSX_ERR_INVALID_ARG;
}
extern "C" bsx_tree *bsx_batch_tree(bsx_batch *b) { return b ? &b->tree : nullptr; }
extern "C" const uint64_t *bsx_batch_lde_device(const bsx_batch *b) { return b ? (const uint64_t *)b->lde : nullptr; }
extern "C" const uint64_t *bsx_batch_digests_device(const bsx_batch *b) {
    return b ? (const uint64_t *)b->tree.digests : nullptr;
}
extern "C" int32_t bsx_batch_shape(const bsx_batch *b, int32_t *log_n, int32_t *n_cols, int32_t *rate_bits,
                                   int32_t *cap_height, int32_t *n_salt) {
    if (!b) return BSX_ERR_INVALID_ARG;
    if (log_n) *log_n = b->log_n;
    if (n_cols) *n_cols = b->n_cols;
    if (rate_bits) *rate_bits = b->rate_bits;
    if (cap_height) *cap_height = b->tree.cap_height;
    if (n_salt) *n_salt = b->n_salt;
    return BSX_OK;
}

static inline u64 host_bitrev(u64 x, int bits) {
    u64 r = 0;
    for (int i = 0; i < bits; i++) r |= ((x >> i) & 1ULL) << (bits - 1 - i);
    return r;
}

// fri/oracle.rs get_lde_values: &merkle_tree.leaves[reverse_bits(index * step, degree_log + rate_bits)]
// [..len - salt_size]
extern "C" int32_t bsx_batch_get_lde_values(bsx_batch *b, uint64_t index, uint64_t step, uint64_t *out_host) {
    if (!b) return BSX_ERR_INVALID_ARG;
    bsx_ctx *ctx = b->ctx;
    BsxScope scope(ctx);
    int log_nb = b->log_n + b->rate_bits;
    u64 i = index * step;
    if (i >> log_nb) return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "index * step out of range");
    if (b->n_cols == 0) return BSX_OK;
    u64 leaf = host_bitrev(i, log_nb);
    CU(ctx, cudaSetDevice(ctx->device));
    u64 *idx_dev, *row_dev;
    TRY(bsx_alloc(ctx, &idx_dev, 1));
    TRY(bsx_alloc(ctx, &row_dev, (size_t)b->n_cols));
    CU(ctx, cudaMemcpyAsync(idx_dev, &leaf, sizeof(u64), cudaMemcpyHostToDevice, ctx->stream));
    launch_gather_rows(b->lde, (size_t)1 << log_nb, b->n_cols, idx_dev, 1, row_dev, ctx->stream);
    CU(ctx, cudaMemcpyAsync(out_host, row_dev, (size_t)b->n_cols * sizeof(u64), cudaMemcpyDeviceToHost, ctx->stream));
    bsx_free(ctx, idx_dev);
    bsx_free(ctx, row_dev);
    CU(ctx, cudaStreamSynchronize(ctx->stream));
    return BSX_OK;
}

// fri/oracle.rs get_lde_values_packed: the rows get_lde_values(index_start + i, step) for i < count (upstream
// transposes them into P::WIDTH-lane packed columns; the ABI returns them row by row, `mem` side).
extern "C" int32_t bsx_batch_get_lde_values_packed(bsx_batch *b, uint64_t index_start, uint64_t count, uint64_t step,
                                                   uint64_t *out, int32_t mem) {
    if (!b) return BSX_ERR_INVALID_ARG;
    bsx_ctx *ctx = b->ctx;
    BsxScope scope(ctx);
    const int log_nb = b->log_n + b->rate_bits;
    if (count == 0 || b->n_cols == 0) return BSX_OK;
    if (!out) return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "output is NULL");
    if (count > ((u64)1 << log_nb) || step > ((u64)1 << log_nb) || ((index_start + count - 1) * step) >> log_nb)
        return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "index * step out of range");
    CU(ctx, cudaSetDevice(ctx->device));
    std::vector<u64> leaf(count);
    for (u64 i = 0; i < count; i++) leaf[i] = host_bitrev((index_start + i) * step, log_nb);
    u64 *idx_dev, *rows_dev = (u64 *)out;
    TRY(bsx_alloc(ctx, &idx_dev, count));
    if (mem != BSX_DEVICE) TRY(bsx_alloc(ctx, &rows_dev, count * (size_t)b->n_cols));
    CU(ctx, cudaMemcpyAsync(idx_dev, leaf.data(), count * sizeof(u64), cudaMemcpyHostToDevice, ctx->stream));
    const u64 max_chunk = 1 << 20;  // rows per gather launch
    for (u64 r = 0; r < count; r += max_chunk) {
        u64 k = count - r < max_chunk ? count - r : max_chunk;
        launch_gather_rows(b->lde, (size_t)1 << log_nb, b->n_cols, idx_dev + r, (int)k, rows_dev + r * b->n_cols,
                           ctx->stream);
    }
    CU(ctx, cudaGetLastError());
    if (mem != BSX_DEVICE) {
        TRY(bsx_copy_out(ctx, (u64 *)out, rows_dev, count * (size_t)b->n_cols, mem));
        bsx_free(ctx, rows_dev);
    }
    bsx_free(ctx, idx_dev);
    CU(ctx, cudaStreamSynchronize(ctx->stream));  // leaf[] is read by the async copy
    return BSX_OK;
}

extern "C" int32_t bsx_batch_coeffs(bsx_batch *b, uint64_t *out, int32_t mem) {
    if (!b || (!out && b->n_cols > 0)) return BSX_ERR_INVALID_ARG;
    bsx_ctx *ctx = b->ctx;
    BsxScope scope(ctx);
    size_t n = (size_t)1 << b->log_n, words = n * (size_t)b->n_cols;
    if (words == 0) return BSX_OK;
    CU(ctx, cudaSetDevice(ctx->device));
    if (b->coeffs_bitrev) {
        u64 *tmp;
        TRY(bsx_alloc(ctx, &tmp, words));
        launch_bitrev_copy(b->coeffs, n, tmp, n, b->log_n, b->n_cols, ctx->stream);
        TRY(bsx_copy_out(ctx, (u64 *)out, tmp, words, mem));
        bsx_free(ctx, tmp);
    } else {
        TRY(bsx_copy_out(ctx, (u64 *)out, b->coeffs, words, mem));
    }
    CU(ctx, cudaStreamSynchronize(ctx->stream));
    return BSX_OK;
}

// Row-major leaves to host memory, pipelined in row slabs: the slab is transposed on the device (LDE viewed as a
// w x nb matrix -> rows), DMA'd into the context's pinned ring on copy_stream, and copied from there into the
// caller's (pageable) buffer by the staging threads while the next slabs are in flight.  Device scratch is two
// slabs, not a second copy of the LDE; a 9 GB export used to need 9 GB of scratch and a driver-staged copy.
static void scatter_copy(u64 *dst, const u64 *src, size_t words, int threads) {
    const size_t piece = (size_t)1 << 17;
    const size_t total = (words + piece - 1) / piece;
    std::atomic<size_t> next(0);
    auto work = [&]() {
        for (;;) {
            size_t t = next.fetch_add(1);
            if (t >= total) return;
            size_t off = t * piece, len = words - off < piece ? words - off : piece;
            memcpy(dst + off, src + off, len * sizeof(u64));
        }
    };
    if ((size_t)threads > total) threads = (int)total;
    std::vector<std::thread> pool;
    try {
        for (int i = 1; i < threads; i++) pool.emplace_back(work);
    } catch (...) {
    }
    work();
    for (auto &th : pool) th.join();
}

static int export_leaves(bsx_batch *b, u64 *leaves_host) {
    bsx_ctx *ctx = b->ctx;
    const size_t nb = (size_t)1 << (b->log_n + b->rate_bits), w = (size_t)(b->n_cols + b->n_salt);
    if (nb == 0 || w == 0) return BSX_OK;
    CU(ctx, cudaSetDevice(ctx->device));
    const bool direct = host_is_pinned(leaves_host);
    size_t rows = ((size_t)4 << 20) / w;  // ~32 MiB slabs
    if (rows < 1) rows = 1;
    if (rows > nb) rows = nb;
    if (!direct) TRY(stage_ring_reserve(ctx, rows * w));
    u64 *dbuf[2];
    TRY(bsx_alloc(ctx, &dbuf[0], rows * w));
    TRY(bsx_alloc(ctx, &dbuf[1], rows * w));
    cudaEvent_t transposed[2], drained[2];
    for (int i = 0; i < 2; i++) {
        cudaEventCreateWithFlags(&transposed[i], cudaEventDisableTiming);
        cudaEventCreateWithFlags(&drained[i], cudaEventDisableTiming);
    }
    const size_t n_slabs = (nb + rows - 1) / rows;
    int rc = BSX_OK;
    // slab i: device buffer i % 2, ring slot i % BSX_STAGE_SLOTS; the host copy of slab i - 2 runs while slab i - 1 is
    // on the bus and slab i is being transposed
    auto host_part = [&](size_t i) {
        const size_t r0 = i * rows, r = nb - r0 < rows ? nb - r0 : rows;
        const int slot = (int)(i % BSX_STAGE_SLOTS);
        if (cudaEventSynchronize(ctx->stage_done[slot]) != cudaSuccess) return bsx_fail(ctx, BSX_ERR_CUDA, "export copy failed");
        scatter_copy(leaves_host + r0 * w, ctx->stage_ring + (size_t)slot * ctx->stage_slot_words, r * w, stage_threads());
        return (int)BSX_OK;
    };
    for (size_t i = 0; i < n_slabs && rc == BSX_OK; i++) {
        const size_t r0 = i * rows, r = nb - r0 < rows ? nb - r0 : rows;
        const int d = (int)(i & 1);
        if (i >= 2) cudaStreamWaitEvent(ctx->stream, drained[d], 0);  // the DMA out of this device buffer is over
        launch_transpose_strided(b->lde + r0, nb, dbuf[d], w, r, ctx->stream);
        cudaEventRecord(transposed[d], ctx->stream);
        cudaStreamWaitEvent(ctx->copy_stream, transposed[d], 0);
        if (direct) {
            cudaMemcpyAsync(leaves_host + r0 * w, dbuf[d], r * w * sizeof(u64), cudaMemcpyDeviceToHost, ctx->copy_stream);
        } else {
            const int slot = (int)(i % BSX_STAGE_SLOTS);
            if (i >= BSX_STAGE_SLOTS) rc = host_part(i - BSX_STAGE_SLOTS);  // frees the slot
            if (rc) break;
            cudaMemcpyAsync(ctx->stage_ring + (size_t)slot * ctx->stage_slot_words, dbuf[d], r * w * sizeof(u64),
                            cudaMemcpyDeviceToHost, ctx->copy_stream);
            cudaEventRecord(ctx->stage_done[slot], ctx->copy_stream);
        }
        cudaEventRecord(drained[d], ctx->copy_stream);
    }
    if (!direct)
        for (size_t i = n_slabs > BSX_STAGE_SLOTS ? n_slabs - BSX_STAGE_SLOTS : 0; i < n_slabs && rc == BSX_OK; i++)
            rc = host_part(i);
    cudaError_t e = cudaStreamSynchronize(ctx->copy_stream);
    if (rc == BSX_OK && (e != cudaSuccess || cudaGetLastError() != cudaSuccess))
        rc = bsx_fail(ctx, BSX_ERR_CUDA, "leaf export failed: %s", cudaGetErrorString(e));
    cudaStreamSynchronize(ctx->stream);
    for (int i = 0; i < 2; i++) {
        cudaEventDestroy(transposed[i]);
        cudaEventDestroy(drained[i]);
    }
    bsx_free(ctx, dbuf[0]);
    bsx_free(ctx, dbuf[1]);
    return rc;
}

extern "C" int32_t bsx_batch_export(bsx_batch *b, uint64_t *coeffs_host, uint64_t *leaves_host,
                                    uint64_t *digests_host) {
    if (!b) return BSX_ERR_INVALID_ARG;
    bsx_ctx *ctx = b->ctx;
    BsxScope scope(ctx);
    if (coeffs_host) TRY(bsx_batch_coeffs(b, coeffs_host, BSX_HOST));
    if (leaves_host) TRY(export_leaves(b, (u64 *)leaves_host));
    if (digests_host) TRY(bsx_merkle_digests(&b->tree, digests_host));
    return BSX_OK;
}

// plonky2's recursive `digests` layout -> level order (inverse of to_plonky2_layout)
static void from_plonky2_layout(u64 *lv, size_t n, int height, size_t node, const u64 *buf, size_t buf_len) {
    if (buf_len == 0) return;
    size_t half = buf_len / 2, off = 0;
    for (int k = 0; k < height - 1; k++) off += n >> k;
    memcpy(lv + 4 * (off + 2 * node), buf + 4 * (half - 1), 4 * sizeof(u64));
    memcpy(lv + 4 * (off + 2 * node + 1), buf + 4 * half, 4 * sizeof(u64));
    from_plonky2_layout(lv, n, height - 1, 2 * node, buf, half - 1);
    from_plonky2_layout(lv, n, height - 1, 2 * node + 1, buf + 4 * (half + 1), half - 1);
}

// A batch restored from its serialised parts (the constants + sigmas commitment stored with the circuit):
// nothing is recomputed, exactly like upstream's reader.
extern "C" int32_t bsx_batch_import(bsx_ctx *ctx, const uint64_t *coeffs_host, const uint64_t *leaves_host,
                                    const uint64_t *digests_host, const uint64_t *cap_host, int32_t log_n,
                                    int32_t n_cols, int32_t n_salt, int32_t rate_bits, int32_t cap_height,
                                    bsx_batch **out) {
    if (!ctx || !out) return BSX_ERR_INVALID_ARG;
    BsxScope scope(ctx);
    *out = nullptr;
    if (log_n < 0 || log_n > 24)
        return bsx_fail(ctx, BSX_ERR_UNSUPPORTED, "degree_log=%d outside the supported range [0, 24]", log_n);
    if (rate_bits < 0 || rate_bits > 8 || log_n + rate_bits > 32)
        return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "bad rate_bits=%d", rate_bits);
    if (n_cols < 0 || n_cols > 65535 || (n_salt != 0 && n_salt != BSX_SALT_SIZE))
        return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "bad n_cols=%d / n_salt=%d", n_cols, n_salt);
    const int log_nb = log_n + rate_bits;
    if (cap_height < 0 || cap_height > log_nb)
        return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "cap_height=%d should be at most log2(leaves.len())=%d",
                        cap_height, log_nb);
    const size_t n = (size_t)1 << log_n, nb = (size_t)1 << log_nb, w = (size_t)(n_cols + n_salt);
    const size_t dig_words = n_digest_words(log_nb, cap_height), cap_words = (size_t)4 << cap_height;
    if ((n_cols > 0 && !coeffs_host) || (w > 0 && !leaves_host) || (dig_words > 0 && !digests_host) || !cap_host)
        return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "a serialised part is NULL");
    CU(ctx, cudaSetDevice(ctx->device));

    bsx_batch *bt = new bsx_batch();
    bt->ctx = ctx;
    bt->log_n = log_n;
    bt->n_cols = n_cols;
    bt->n_salt = n_salt;
    bt->rate_bits = rate_bits;
    bt->coeffs = bt->lde = nullptr;
    bt->coeffs_bitrev = false;
    bt->tree.ctx = ctx;
    bt->tree.leaves = nullptr;
    bt->tree.digests = bt->tree.cap_dev = nullptr;
    bt->tree.owns_leaves = false;
    bt->tree.embedded = true;
    bt->tree.log_leaves = log_nb;
    bt->tree.leaf_len = (int)w;
    bt->tree.cap_height = cap_height;
    bt->tree.col_stride = nb;
    u64 *tmp = nullptr;
    int rc = BSX_OK;
    do {
        if ((rc = bsx_alloc_owned(ctx, &bt->coeffs, n * n_cols))) break;
        if ((rc = bsx_alloc_owned(ctx, &bt->lde, nb * w))) break;
        if ((rc = bsx_tree_alloc(&bt->tree))) break;
        bt->tree.leaves = bt->lde;
        if ((rc = bsx_copy_in(ctx, bt->coeffs, (const u64 *)coeffs_host, n * n_cols, BSX_HOST))) break;
        if (w > 0) {
            if ((rc = bsx_alloc(ctx, &tmp, nb * w))) break;
            if ((rc = bsx_copy_in(ctx, tmp, (const u64 *)leaves_host, nb * w, BSX_HOST))) break;
            launch_transpose(tmp, bt->lde, nb, w, ctx->stream);  // (nb x w) row-major leaves -> column-major LDE
        }
        std::vector<u64> lv(dig_words);
        if (dig_words > 0) {
            const int height = log_nb - cap_height;
            const size_t sub = ((size_t)2 << height) - 2;
            for (size_t c = 0; c < ((size_t)1 << cap_height); c++)
                from_plonky2_layout(lv.data(), nb, height, c, (const u64 *)digests_host + 4 * c * sub, sub);
            if ((rc = bsx_copy_in(ctx, bt->tree.digests, lv.data(), dig_words, BSX_HOST))) break;
        }
        bt->tree.cap_host.assign((const u64 *)cap_host, (const u64 *)cap_host + cap_words);
        if ((rc = bsx_copy_in(ctx, bt->tree.cap_dev, bt->tree.cap_host.data(), cap_words, BSX_HOST))) break;
        cudaError_t e = cudaStreamSynchronize(ctx->stream);  // lv and the caller's buffers may go away now
        if (e == cudaSuccess) e = cudaGetLastError();
        if (e != cudaSuccess) rc = bsx_fail(ctx, BSX_ERR_CUDA, "import failed: %s", cudaGetErrorString(e));
    } while (0);
    bsx_free(ctx, tmp);
    if (rc != BSX_OK) {
        bsx_tree_release(&bt->tree);
        bsx_free(ctx, bt->coeffs);
        bsx_free(ctx, bt->lde);
        delete bt;
        return rc;
    }
    *out = bt;
    return BSX_OK;
}

// ------------------------------------------------------------------------------------ test hooks
extern "C" int32_t bsx_poseidon_permute(bsx_ctx *ctx, uint64_t *states, int32_t mem, size_t n) {
    if (!ctx || (!states && n)) return BSX_ERR_INVALID_ARG;
    BsxScope scope(ctx);
    if (n == 0) return BSX_OK;
    CU(ctx, cudaSetDevice(ctx->device));
    u64 *d = (u64 *)states;
    if (mem != BSX_DEVICE) {
        TRY(bsx_alloc(ctx, &d, n * 12));
        TRY(bsx_copy_in(ctx, d, (const u64 *)states, n * 12, mem));
    }
    launch_poseidon_permute(d, n, ctx->stream);
    CU(ctx, cudaGetLastError());
    if (mem != BSX_DEVICE) {
        TRY(bsx_copy_out(ctx, (u64 *)states, d, n * 12, mem));
        bsx_free(ctx, d);
    }
    CU(ctx, cudaStreamSynchronize(ctx->stream));
    return BSX_OK;
}

// test hook: n words of the ChaCha20 key stream the blinding salts are drawn from (raw: no rejection mod p)
extern "C" int32_t bsx_debug_salt_stream(bsx_ctx *ctx, const uint32_t key[8], int32_t raw, uint64_t *out_host, size_t n) {
    if (!ctx || !key || (!out_host && n)) return BSX_ERR_INVALID_ARG;
    BsxScope scope(ctx);
    if (n == 0) return BSX_OK;
    CU(ctx, cudaSetDevice(ctx->device));
    u64 *d;
    TRY(bsx_alloc(ctx, &d, n));
    launch_fill_random(d, n, key, raw, ctx->stream);
    TRY(bsx_copy_out(ctx, (u64 *)out_host, d, n, BSX_HOST));
    bsx_free(ctx, d);
    CU(ctx, cudaStreamSynchronize(ctx->stream));
    return BSX_OK;
}

// values (natural order) -> coefficients (natural order) of the polynomial that takes them on shift * H:
// PolynomialValues::coset_ifft = ifft, then c_i *= shift^-i.  In place on the device buffer `buf`; `tmp` is scratch.
static int coset_or_plain_ntt(bsx_ctx *ctx, u64 *buf, u64 *tmp, int log_n, int n_cols, int kind, u64 shift) {
    size_t n = (size_t)1 << log_n;
    NttPlan plan;
    plan.n_dest = 0;
    plan.log_n = log_n;
    bsx_split(log_n, &plan.a, &plan.b);
    plan.inverse = kind == 1 || kind == 3;
    plan.src_bitrev = 0;
    plan.n_cosets = 1;
    plan.pre_hi = plan.pre_lo = nullptr;
    if (kind == 2) TRY(bsx_coset_tables(ctx, log_n, 0, shift % GL_P, &plan.pre_hi, &plan.pre_lo));
    TRY(bsx_pow_table(ctx, log_n, plan.inverse, &plan.tw));
    plan.scale = plan.inverse ? gl_host_inv((u64)n % GL_P) : 1;
    TRY(bsx_post_table(ctx, &plan, kind == 2 ? shift % GL_P : 0, n_cols));
    launch_ntt(ctx->tables, plan, buf, n, buf, n, n_cols, ctx->stream);
    launch_bitrev_copy(buf, n, tmp, n, log_n, n_cols, ctx->stream);
    if (kind == 3) {
        const u64 *hi, *lo;
        TRY(bsx_coset_tables(ctx, log_n, 0, gl_host_inv(shift % GL_P), &hi, &lo));
        launch_scale_two_level(tmp, n, log_n, plan.b, n_cols, hi, lo, ctx->stream);
    }
    CU(ctx, cudaGetLastError());
    return BSX_OK;
}

extern "C" int32_t bsx_ntt(bsx_ctx *ctx, uint64_t *data, int32_t mem, int32_t log_n, int32_t n_cols, int32_t kind,
                           uint64_t shift) {
    if (!ctx || (!data && n_cols > 0)) return BSX_ERR_INVALID_ARG;
    BsxScope scope(ctx);
    if (log_n < 0 || log_n > 24) return bsx_fail(ctx, BSX_ERR_UNSUPPORTED, "log_n=%d outside [0, 24]", log_n);
    if (kind < 0 || kind > 3 || n_cols < 0 || n_cols > 65535) return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "bad kind/n_cols");
    if (kind == 3 && shift % GL_P == 0) return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "coset shift must be non-zero");
    if (n_cols == 0) return BSX_OK;
    CU(ctx, cudaSetDevice(ctx->device));
    size_t n = (size_t)1 << log_n, words = n * (size_t)n_cols;
    u64 *buf, *tmp;
    TRY(bsx_alloc(ctx, &buf, words));
    TRY(bsx_alloc(ctx, &tmp, words));
    TRY(bsx_copy_in(ctx, buf, (const u64 *)data, words, mem));
    TRY(coset_or_plain_ntt(ctx, buf, tmp, log_n, n_cols, kind, shift));
    TRY(bsx_copy_out(ctx, (u64 *)data, tmp, words, mem));
    bsx_free(ctx, buf);
    bsx_free(ctx, tmp);
    CU(ctx, cudaStreamSynchronize(ctx->stream));
    return BSX_OK;
}

// plonk/prover.rs: the tail of compute_quotient_polys (`values.coset_ifft(F::coset_shift())` per challenge) and the
// commit that follows it in prove() (`quotient_poly.trim_to_len(quotient_degree)`, `.chunks(degree)`,
// `PolynomialBatch::from_coeffs(all_quotient_poly_chunks, ...)`), without the coefficients leaving the device.
extern "C" int32_t bsx_batch_from_quotient_values(bsx_ctx *ctx, const uint64_t *values, int32_t mem, int32_t degree_log,
                                                  int32_t quotient_degree_bits, int32_t quotient_degree_factor,
                                                  int32_t n_polys, int32_t rate_bits, int32_t cap_height,
                                                  int32_t blinding, const uint64_t *salt, bsx_batch **out,
                                                  bsx_timing *timing) {
    if (!ctx || !out || (!values && n_polys > 0)) return BSX_ERR_INVALID_ARG;
    *out = nullptr;
    BsxScope scope(ctx);
    const int log_big = degree_log + quotient_degree_bits;
    if (degree_log < 0 || quotient_degree_bits < 0 || log_big > 24)
        return bsx_fail(ctx, BSX_ERR_UNSUPPORTED, "degree_log=%d + quotient_degree_bits=%d outside [0, 24]", degree_log,
                        quotient_degree_bits);
    if (quotient_degree_factor < 1 || quotient_degree_factor > (1 << quotient_degree_bits))
        return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "quotient_degree_factor=%d does not fit 2^%d", quotient_degree_factor,
                        quotient_degree_bits);
    if (n_polys < 1 || n_polys * quotient_degree_factor > 65535) return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "bad n_polys");
    CU(ctx, cudaSetDevice(ctx->device));
    const size_t n = (size_t)1 << degree_log, big = (size_t)1 << log_big, words = big * (size_t)n_polys;
    u64 *buf, *tmp;
    int *flag;
    TRY(bsx_alloc(ctx, &buf, words));
    TRY(bsx_alloc(ctx, &tmp, words));
    TRY(bsx_alloc(ctx, (u64 **)&flag, 1));
    TRY(bsx_copy_in(ctx, buf, (const u64 *)values, words, mem));
    TRY(coset_or_plain_ntt(ctx, buf, tmp, log_big, n_polys, 3, 7));  // F::coset_shift() = 7
    const size_t keep = (size_t)quotient_degree_factor * n;
    if (keep < big) {
        // trim_to_len: the dropped coefficients must all be zero ("Quotient has failed, the vanishing polynomial
        // is not divisible by Z_H" upstream); the kept chunks are packed to factor * N words per polynomial
        CU(ctx, cudaMemsetAsync(flag, 0, sizeof(u64), ctx->stream));
        launch_any_nonzero(tmp, big, keep, big, n_polys, flag, ctx->stream);
        int host_flag = 0;
        CU(ctx, cudaMemcpyAsync(&host_flag, flag, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
        CU(ctx, cudaStreamSynchronize(ctx->stream));
        if (host_flag)
            return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "quotient polynomial has degree >= %d * 2^%d (trim_to_len fails upstream)",
                            quotient_degree_factor, degree_log);
        // compaction into the other buffer (`buf` is free after the transform): in place the ranges of
        // polynomial c overlap whenever keep > big / 2, which cudaMemcpy does not allow
        CU(ctx, cudaMemcpy2DAsync(buf, keep * sizeof(u64), tmp, big * sizeof(u64), keep * sizeof(u64), (size_t)n_polys,
                                  cudaMemcpyDeviceToDevice, ctx->stream));
        std::swap(buf, tmp);
    }
    // `salt` lives on the `mem` side like `values`; the commit below takes device pointers
    u64 *salt_dev = nullptr;
    if (blinding && salt && mem != BSX_DEVICE) {
        const size_t sw = ((size_t)BSX_SALT_SIZE << (degree_log + rate_bits));
        TRY(bsx_alloc(ctx, &salt_dev, sw));
        TRY(bsx_copy_in(ctx, salt_dev, (const u64 *)salt, sw, mem));
        salt = (const uint64_t *)salt_dev;
    }
    // polynomial c, chunk j = words [(c * factor + j) * N, +N): exactly a column-major (n_polys * factor) x N matrix
    int32_t rc = bsx_batch_from_coeffs(ctx, (const uint64_t *)tmp, BSX_DEVICE, degree_log, n_polys * quotient_degree_factor, rate_bits,
                                       cap_height, blinding, salt, out, timing);
    bsx_free(ctx, salt_dev);
    bsx_free(ctx, buf);
    bsx_free(ctx, tmp);
    bsx_free(ctx, (u64 *)flag);
    return rc;
}
