// api_shard.cu -- one PolynomialBatch commit sharded over the GPUs of a box.
//
// plonky2 has no multi-device path; this is the partitioning BASELINE.json's north star lays out for
// `PolynomialBatch::from_values` (fri/oracle.rs):
//   1. rank g owns the columns [c0_g, c1_g): iNTT + coset LDE per column need no communication;
//   2. rank g owns the leaf rows [g * NB/G, (g+1) * NB/G) of EVERY column.  The DIF transforms leave each
//      column in leaf order, so that slab is a contiguous range of every column -- no strided gather;
//   3. the last NTT pass stores each output word straight into the owning rank's slab (`dests` of
//      bsx_commit_transforms).  With peer pointers (CUDA IPC across processes, or peer access inside one
//      process) the exchange is NVLink P2P stores issued by the transform kernel itself, overlapped with
//      its arithmetic; with a staging buffer the slabs are laid out destination-major for one
//      NCCL all-to-all issued by the host (torch.distributed / ncclSend+ncclRecv);
//   4. each rank hashes its rows and builds its 2^cap_height / G cap subtrees; the host gathers the caps.
// Every byte of the result is identical to the single-GPU batch: same LDE words, same digests, same cap.
#include "api_internal.h"

#include <cstring>
#include <vector>

struct bsx_shard {
    bsx_ctx *ctx;
    int log_n, n_cols_total, rate_bits, cap_height, rank, n_ranks, log_ranks;
    int c0, c1;            // this rank's columns
    size_t chunk;          // leaves per rank = 2^(log_n + rate_bits) / n_ranks
    u64 *slab;             // n_cols_total x chunk, column-major: this rank's leaf rows (cudaMalloc: IPC-able)
    u64 *peer[BSX_MAX_RANKS];   // every rank's slab as seen from this device (peer[rank] == slab)
    bool peer_ipc[BSX_MAX_RANKS];
    bool peer_same_device[BSX_MAX_RANKS];  // an IPC peer that lives on this rank's own GPU (see device_tag)
    u64 device_tag;
    u64 *staging;          // optional destination-major send buffer: n_ranks x (c1 - c0) x chunk
    u64 *coeffs;           // (c1 - c0) x 2^log_n
    bool coeffs_bitrev;
    // Mailbox behind the slab (same allocation, so a peer that mapped the slab sees it): G arrival flags, G cap flags,
    // G local caps.  Written by the peers' barrier kernels (bsx_shard_run_p2p), epochs only grow.
    size_t slab_words, cap_words;
    u64 epoch;
    int *timeout_flag;     // pinned host word the barrier kernel raises when a peer never arrives
};
static inline u64 *shard_mail(const bsx_shard *s, int rank) { return s->peer[rank] ? s->peer[rank] + s->slab_words : nullptr; }

extern "C" int32_t bsx_shard_column_range(int32_t n_cols_total, int32_t rank, int32_t n_ranks, int32_t *c0,
                                          int32_t *c1) {
    if (n_cols_total < 0 || n_ranks <= 0 || rank < 0 || rank >= n_ranks || !c0 || !c1) return BSX_ERR_INVALID_ARG;
    // balanced contiguous ranges: the first (C mod G) ranks take one extra column (135 = 7 * 17 + 16)
    int q = n_cols_total / n_ranks, r = n_cols_total % n_ranks;
    *c0 = rank * q + (rank < r ? rank : r);
    *c1 = *c0 + q + (rank < r ? 1 : 0);
    return BSX_OK;
}

extern "C" int32_t bsx_shard_create(bsx_ctx *ctx, int32_t log_n, int32_t n_cols_total, int32_t rate_bits,
                                    int32_t cap_height, int32_t rank, int32_t n_ranks, bsx_shard **out) {
    if (!ctx || !out) return BSX_ERR_INVALID_ARG;
    *out = nullptr;
    if (log_n < 0 || log_n > 24) return bsx_fail(ctx, BSX_ERR_UNSUPPORTED, "degree_log=%d outside [0, 24]", log_n);
    if (rate_bits < 0 || rate_bits > 8) return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "bad rate_bits=%d", rate_bits);
    if (n_cols_total < 0 || n_cols_total > 65535) return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "bad n_cols=%d", n_cols_total);
    int log_ranks = 0;
    while ((1 << log_ranks) < n_ranks) log_ranks++;
    if (n_ranks < 1 || n_ranks > BSX_MAX_RANKS || (1 << log_ranks) != n_ranks)
        return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "n_ranks=%d must be a power of two <= %d", n_ranks, BSX_MAX_RANKS);
    if (rank < 0 || rank >= n_ranks) return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "rank %d of %d", rank, n_ranks);
    const int log_nb = log_n + rate_bits;
    if (cap_height < 0 || cap_height > log_nb)
        return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "cap_height=%d should be at most log2(leaves.len())=%d", cap_height,
                        log_nb);
    if (log_ranks > log_nb) return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "more ranks than leaves");
    CU(ctx, cudaSetDevice(ctx->device));
    bsx_shard *s = new bsx_shard();
    s->ctx = ctx;
    s->log_n = log_n;
    s->n_cols_total = n_cols_total;
    s->rate_bits = rate_bits;
    s->cap_height = cap_height;
    s->rank = rank;
    s->n_ranks = n_ranks;
    s->log_ranks = log_ranks;
    bsx_shard_column_range(n_cols_total, rank, n_ranks, &s->c0, &s->c1);
    s->chunk = (size_t)1 << (log_nb - log_ranks);
    s->slab = s->staging = s->coeffs = nullptr;
    s->coeffs_bitrev = false;
    for (int i = 0; i < BSX_MAX_RANKS; i++) {
        s->peer[i] = nullptr;
        s->peer_ipc[i] = false;
        s->peer_same_device[i] = false;
    }
    s->slab_words = s->chunk * (size_t)n_cols_total;
    s->cap_words = (size_t)4 << (cap_height > log_ranks ? cap_height - log_ranks : 0);
    s->epoch = 0;
    s->timeout_flag = nullptr;
    // + one word behind the mailbox: a tag of the physical device (PCI domain : bus : device), so that a peer mapping the
    // slab through IPC can tell whether it shares this GPU
    const size_t mail_words = (size_t)n_ranks * (2 + s->cap_words) + 1;
    {
        cudaDeviceProp prop;
        s->device_tag = 0;
        if (cudaGetDeviceProperties(&prop, ctx->device) == cudaSuccess)
            s->device_tag = ((u64)1 << 63) | ((u64)(unsigned)prop.pciDomainID << 32) | ((u64)(unsigned)prop.pciBusID << 16) |
                            (u64)(unsigned)prop.pciDeviceID;
        cudaGetLastError();
    }
    size_t bytes = (s->slab_words + mail_words) * sizeof(u64);
    if (bytes) {
        cudaError_t e = cudaMalloc((void **)&s->slab, bytes);  // not the async pool: cudaIpcGetMemHandle needs cudaMalloc
        if (e != cudaSuccess) {
            cudaGetLastError();
            delete s;
            return bsx_fail(ctx, BSX_ERR_OOM, "row slab of %zu bytes: %s", bytes, cudaGetErrorString(e));
        }
        cudaMemsetAsync(s->slab + s->slab_words, 0, mail_words * sizeof(u64), ctx->stream);
        cudaMemcpyAsync(s->slab + s->slab_words + mail_words - 1, &s->device_tag, sizeof(u64), cudaMemcpyHostToDevice, ctx->stream);
        cudaStreamSynchronize(ctx->stream);  // before any peer can learn the address
    }
    s->peer[rank] = s->slab;
    *out = s;
    return BSX_OK;
}

extern "C" void bsx_shard_free(bsx_shard *s) {
    if (!s) return;
    cudaSetDevice(s->ctx->device);
    cudaStreamSynchronize(s->ctx->stream);
    for (int i = 0; i < s->n_ranks; i++)
        if (s->peer_ipc[i] && s->peer[i]) cudaIpcCloseMemHandle(s->peer[i]);
    if (s->slab) cudaFree(s->slab);
    if (s->timeout_flag) cudaFreeHost(s->timeout_flag);
    bsx_free(s->ctx, s->staging);
    bsx_free(s->ctx, s->coeffs);
    delete s;
}

extern "C" uint64_t *bsx_shard_slab_device(bsx_shard *s) { return s ? (uint64_t *)s->slab : nullptr; }
extern "C" uint64_t *bsx_shard_staging_device(bsx_shard *s) { return s ? (uint64_t *)s->staging : nullptr; }
extern "C" const uint64_t *bsx_shard_coeffs_device(const bsx_shard *s) { return s ? (const uint64_t *)s->coeffs : nullptr; }

extern "C" int32_t bsx_shard_slab_ipc(bsx_shard *s, uint8_t *handle_out) {
    if (!s || !handle_out) return BSX_ERR_INVALID_ARG;
    static_assert(sizeof(cudaIpcMemHandle_t) == BSX_IPC_HANDLE_BYTES, "IPC handle size");
    memset(handle_out, 0, BSX_IPC_HANDLE_BYTES);
    if (!s->slab) return BSX_OK;
    CU(s->ctx, cudaSetDevice(s->ctx->device));
    cudaIpcMemHandle_t h;
    CU(s->ctx, cudaIpcGetMemHandle(&h, s->slab));
    memcpy(handle_out, &h, sizeof(h));
    return BSX_OK;
}

extern "C" int32_t bsx_shard_set_peer_ipc(bsx_shard *s, int32_t peer_rank, const uint8_t *handle) {
    if (!s || !handle || peer_rank < 0 || peer_rank >= s->n_ranks) return BSX_ERR_INVALID_ARG;
    if (peer_rank == s->rank || !s->slab) return BSX_OK;
    CU(s->ctx, cudaSetDevice(s->ctx->device));
    cudaIpcMemHandle_t h;
    memcpy(&h, handle, sizeof(h));
    void *p = nullptr;
    CU(s->ctx, cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    s->peer[peer_rank] = (u64 *)p;
    s->peer_ipc[peer_rank] = true;
    {   // does the peer live on this very GPU?  (two processes on one device: the device-side barrier must not be used)
        u64 tag = 0;
        const size_t tag_at = s->slab_words + (size_t)s->n_ranks * (2 + s->cap_words);
        CU(s->ctx, cudaMemcpy(&tag, (u64 *)p + tag_at, sizeof(u64), cudaMemcpyDeviceToHost));
        s->peer_same_device[peer_rank] = tag != 0 && tag == s->device_tag;
    }
    return BSX_OK;
}

extern "C" int32_t bsx_shard_set_peer_ptr(bsx_shard *s, int32_t peer_rank, uint64_t *slab_dev) {
    if (!s || peer_rank < 0 || peer_rank >= s->n_ranks) return BSX_ERR_INVALID_ARG;
    if (peer_rank == s->rank) return BSX_OK;
    CU(s->ctx, cudaSetDevice(s->ctx->device));
    if (slab_dev) {
        cudaPointerAttributes at;
        CU(s->ctx, cudaPointerGetAttributes(&at, slab_dev));
        if (at.type == cudaMemoryTypeDevice && at.device != s->ctx->device) {
            cudaError_t e = cudaDeviceEnablePeerAccess(at.device, 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) {
                cudaGetLastError();
                return bsx_fail(s->ctx, BSX_ERR_CUDA, "no peer access from device %d to %d: %s", s->ctx->device,
                                at.device, cudaGetErrorString(e));
            }
            cudaGetLastError();
        }
    }
    s->peer[peer_rank] = (u64 *)slab_dev;
    s->peer_ipc[peer_rank] = false;
    return BSX_OK;
}

// Transforms of the local columns.  use_staging == 0: every row slab must be known (bsx_shard_set_peer_*);
// the LDE words go straight to their owners.  use_staging != 0: they go to a local destination-major
// buffer, block d = this rank's columns x rank d's rows, for the host's all-to-all; block d must land at
// slab_d + c0 * chunk on rank d.
// Enqueues copy-in and transforms on ctx->stream (a pinned host input is uploaded in column chunks on the copy stream,
// chunk k + 1 under the transforms of chunk k); no host synchronisation.  ev: optional [after copy-in of the first
// chunk, after the last iNTT, after the last LDE].
static int shard_transform_enqueue(bsx_shard *s, const uint64_t *input, int32_t mem, int32_t input_is_coeffs,
                                   int32_t use_staging, cudaEvent_t *ev, float *ifft_ms_acc, std::vector<cudaEvent_t> *chunk_ev) {
    bsx_ctx *ctx = s->ctx;
    const int n_local = s->c1 - s->c0;
    if (!input && n_local > 0) return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "input is NULL");
    const size_t n = (size_t)1 << s->log_n, nb = n << s->rate_bits;
    BsxDests D;
    D.n = s->n_ranks;
    D.lc = s->log_n + s->rate_bits - s->log_ranks;
    if (use_staging) {
        if (!s->staging) TRY(bsx_alloc_owned(ctx, &s->staging, nb * (size_t)n_local));
        for (int d = 0; d < s->n_ranks; d++) D.ptr[d] = s->staging + (size_t)d * s->chunk * n_local;
    } else {
        for (int d = 0; d < s->n_ranks; d++) {
            if (!s->peer[d] && s->chunk != 0 && s->n_cols_total != 0)
                return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "row slab of rank %d is not set", d);
            D.ptr[d] = s->peer[d] + (size_t)s->c0 * s->chunk;
        }
    }
    bsx_free(ctx, s->coeffs);
    s->coeffs = nullptr;
    (void)ifft_ms_acc;
    if (n_local == 0) return BSX_OK;
    TRY(bsx_alloc_owned(ctx, &s->coeffs, n * n_local));
    // the two-pass transforms keep their intermediate in a local column-major scratch; single-pass
    // sizes (log_n <= 12) write their only pass straight to the destinations
    u64 *scratch = nullptr;
    TRY(bsx_alloc(ctx, &scratch, nb * (size_t)n_local));
    cudaPointerAttributes at;
    const bool pinned = mem == BSX_HOST && cudaPointerGetAttributes(&at, input) == cudaSuccess && at.type == cudaMemoryTypeHost;
    cudaGetLastError();
    int rc = BSX_OK;
    if (pinned && n_local >= 8 && s->log_n >= 13) {
        // Only the first chunk's copy is exposed, so the schedule starts small; chunk sizes follow the transform grid
        // (128 tiles per column at 2^20 rows against 444 resident CTAs: 3 or 7 columns fill whole waves, 4 would leave
        // the second wave 15 % full): 3, 3, 4, then about three chunks of a multiple of seven.
        std::vector<int> cstart;
        {
            int c = 0;
            for (int r : {3, 3, 4})
                if (n_local - c >= r + 3) {
                    cstart.push_back(c);
                    c += r;
                }
            // the rest in about three chunks, each a multiple of seven columns
            int per = ((n_local - c + 2) / 3 + 6) / 7 * 7;
            if (per < 7) per = 7;
            while (c < n_local) {
                cstart.push_back(c);
                c += (n_local - c >= per + 3) ? per : n_local - c;
            }
            cstart.push_back(n_local);
        }
        const int n_chunks = (int)cstart.size() - 1;
        std::vector<cudaEvent_t> &e = *chunk_ev;
        e.resize((size_t)n_chunks + 1);
        for (auto &x : e) cudaEventCreateWithFlags(&x, cudaEventDisableTiming);
        cudaEventRecord(e[(size_t)n_chunks], ctx->stream);  // the buffers exist (stream-ordered allocation)
        cudaStreamWaitEvent(ctx->copy_stream, e[(size_t)n_chunks], 0);
        for (int k = 0; k < n_chunks; k++) {
            const int c0 = cstart[k], cc = cstart[k + 1] - c0;
            CU(ctx, cudaMemcpyAsync(s->coeffs + (size_t)c0 * n, (const u64 *)input + (size_t)c0 * n, (size_t)cc * n * sizeof(u64),
                                    cudaMemcpyHostToDevice, ctx->copy_stream));
            cudaEventRecord(e[(size_t)k], ctx->copy_stream);
        }
        for (int k = 0; k < n_chunks && rc == BSX_OK; k++) {
            const int c0 = cstart[k], cc = cstart[k + 1] - c0;
            cudaStreamWaitEvent(ctx->stream, e[(size_t)k], 0);
            if (k == 0 && ev) cudaEventRecord(ev[0], ctx->stream);
            BsxDests Dk = D;
            for (int d = 0; d < D.n; d++) Dk.ptr[d] = D.ptr[d] + (size_t)c0 * s->chunk;
            rc = bsx_commit_transforms(ctx, s->coeffs + (size_t)c0 * n, s->log_n, cc, s->rate_bits, input_is_coeffs != 0,
                                       scratch + (size_t)c0 * nb, nb, &Dk, nullptr);
        }
        if (ev) cudaEventRecord(ev[1], ctx->stream);  // chunked: iNTT and LDE interleave, reported together as lde
    } else {
        TRY(bsx_copy_in(ctx, s->coeffs, (const u64 *)input, n * n_local, mem));
        if (ev) cudaEventRecord(ev[0], ctx->stream);
        rc = bsx_commit_transforms(ctx, s->coeffs, s->log_n, n_local, s->rate_bits, input_is_coeffs != 0, scratch, nb, &D,
                                   ev ? ev[1] : nullptr);
    }
    if (rc == BSX_OK) s->coeffs_bitrev = !input_is_coeffs;
    if (ev) cudaEventRecord(ev[2], ctx->stream);
    bsx_free(ctx, scratch);
    return rc;
}

// Transforms of the local columns.  use_staging == 0: every row slab must be known (bsx_shard_set_peer_*);
// the LDE words go straight to their owners.  use_staging != 0: they go to a local destination-major
// buffer, block d = this rank's columns x rank d's rows, for the host's all-to-all; block d must land at
// slab_d + c0 * chunk on rank d.
extern "C" int32_t bsx_shard_transform(bsx_shard *s, const uint64_t *input, int32_t mem, int32_t input_is_coeffs,
                                       int32_t use_staging, bsx_timing *timing) {
    if (!s) return BSX_ERR_INVALID_ARG;
    bsx_ctx *ctx = s->ctx;
    BsxScope scope(ctx);
    CU(ctx, cudaSetDevice(ctx->device));
    cudaEvent_t ev[3] = {nullptr, nullptr, nullptr};
    if (timing)
        for (auto &e : ev) cudaEventCreate(&e);
    std::vector<cudaEvent_t> chunk_ev;
    int rc = shard_transform_enqueue(s, input, mem, input_is_coeffs, use_staging, timing ? ev : nullptr, nullptr, &chunk_ev);
    cudaError_t e = cudaStreamSynchronize(ctx->stream);  // stores to peer slabs are complete and visible after this
    if (rc == BSX_OK && e != cudaSuccess) rc = bsx_fail(ctx, BSX_ERR_CUDA, "sharded transform failed: %s", cudaGetErrorString(e));
    for (auto &x : chunk_ev) cudaEventDestroy(x);
    if (timing) {
        memset(timing, 0, sizeof(*timing));
        if (rc == BSX_OK && s->c1 > s->c0) {
            cudaEventElapsedTime(&timing->ifft_ms, ev[0], ev[1]);
            cudaEventElapsedTime(&timing->lde_ms, ev[1], ev[2]);
            timing->total_ms = timing->ifft_ms + timing->lde_ms;
        }
        for (auto &x : ev) cudaEventDestroy(x);
    }
    return rc;
}

// The transforms of the local columns [col_begin, col_begin + col_count) only (input: those columns), so that a host
// running the staged exchange can send the finished columns of one chunk while the next chunk is being transformed.
// Call with increasing col_begin starting at 0; sync == 0 returns with the work enqueued on the context's stream.
extern "C" int32_t bsx_shard_transform_cols(bsx_shard *s, const uint64_t *input, int32_t mem, int32_t input_is_coeffs,
                                            int32_t use_staging, int32_t col_begin, int32_t col_count, int32_t sync) {
    if (!s) return BSX_ERR_INVALID_ARG;
    bsx_ctx *ctx = s->ctx;
    BsxScope scope(ctx);
    const int n_local = s->c1 - s->c0;
    if (col_begin < 0 || col_count < 0 || col_begin + col_count > n_local)
        return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "columns [%d, %d) outside this rank's %d columns", col_begin, col_begin + col_count, n_local);
    if (!input && col_count > 0) return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "input is NULL");
    CU(ctx, cudaSetDevice(ctx->device));
    const size_t n = (size_t)1 << s->log_n, nb = n << s->rate_bits;
    if (col_begin == 0) {
        bsx_free(ctx, s->coeffs);
        s->coeffs = nullptr;
        if (n_local > 0) TRY(bsx_alloc_owned(ctx, &s->coeffs, n * n_local));
    }
    if (col_count > 0) {
        if (!s->coeffs) return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "start with col_begin = 0");
        BsxDests D;
        D.n = s->n_ranks;
        D.lc = s->log_n + s->rate_bits - s->log_ranks;
        if (use_staging) {
            if (!s->staging) TRY(bsx_alloc_owned(ctx, &s->staging, nb * (size_t)n_local));
            for (int d = 0; d < s->n_ranks; d++) D.ptr[d] = s->staging + (size_t)d * s->chunk * n_local + (size_t)col_begin * s->chunk;
        } else {
            for (int d = 0; d < s->n_ranks; d++) {
                if (!s->peer[d]) return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "row slab of rank %d is not set", d);
                D.ptr[d] = s->peer[d] + (size_t)(s->c0 + col_begin) * s->chunk;
            }
        }
        u64 *cf = s->coeffs + (size_t)col_begin * n, *scratch;
        TRY(bsx_copy_in(ctx, cf, (const u64 *)input, n * (size_t)col_count, mem));
        TRY(bsx_alloc(ctx, &scratch, nb * (size_t)col_count));
        TRY(bsx_commit_transforms(ctx, cf, s->log_n, col_count, s->rate_bits, input_is_coeffs != 0, scratch, nb, &D, nullptr));
        bsx_free(ctx, scratch);
        s->coeffs_bitrev = !input_is_coeffs;
    }
    if (sync) CU(ctx, cudaStreamSynchronize(ctx->stream));
    return BSX_OK;
}

// ---- device-side barrier over the mailboxes (one rank per GPU) ------------------------------------------------
// Thread p tells rank p "rank `rank` has reached `epoch`" (slot 0: its stores into your slab are complete -- they were
// issued by kernels that finished earlier on this stream; slot 1: its local cap is in your mailbox), then waits for
// rank p's word in this rank's own mailbox.  Ranks run on different GPUs, so every peer's kernel is resident or will
// be launched regardless of this one spinning; a peer that never arrives (a crashed process) raises *timeout instead
// of hanging the GPU.
struct ShardMail {
    u64 *mail[BSX_MAX_RANKS];
};
__global__ void shard_barrier_kernel(ShardMail M, int rank, int n_ranks, u64 epoch, int slot, const u64 *cap_src,
                                     int cap_words, int *timeout) {
    const int p = threadIdx.x;
    if (p >= n_ranks) return;
    if (cap_src) {
        u64 *dst = M.mail[p] + 2 * n_ranks + (size_t)rank * cap_words;
        for (int i = 0; i < cap_words; i++) dst[i] = cap_src[i];
    }
    __threadfence_system();
    *((volatile u64 *)(M.mail[p] + slot * n_ranks + rank)) = epoch;
    volatile u64 *mine = (volatile u64 *)(M.mail[rank] + slot * n_ranks + p);
    const long long t0 = clock64();
    while (*mine < epoch) {
        if (clock64() - t0 > 20000000000LL) {  // ~10 s
            *timeout = 1;
            break;
        }
        __nanosleep(200);
    }
    __threadfence_system();
}

// One sharded commit with NO host synchronisation between its phases (one process per GPU, peers mapped through
// bsx_shard_set_peer_ipc): transforms whose last pass stores into the owners' slabs -> device barrier -> leaf hashing
// and cap subtrees -> every rank's local cap published into every mailbox -> rank_caps_host (n_ranks x local cap,
// rank order).  The second barrier also makes the slabs safe to overwrite by the next call.
extern "C" int32_t bsx_shard_run_p2p(bsx_shard *s, const uint64_t *input, int32_t mem, int32_t input_is_coeffs,
                                     bsx_tree **tree_out, uint64_t *rank_caps_host, bsx_timing *timing) {
    if (!s || !tree_out || !rank_caps_host) return BSX_ERR_INVALID_ARG;
    bsx_ctx *ctx = s->ctx;
    BsxScope scope(ctx);
    *tree_out = nullptr;
    CU(ctx, cudaSetDevice(ctx->device));
    ShardMail M;
    for (int p = 0; p < BSX_MAX_RANKS; p++) M.mail[p] = nullptr;
    for (int p = 0; p < s->n_ranks; p++) {
        if (!s->peer[p]) return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "row slab of rank %d is not set", p);
        if (p != s->rank) {
            bool same = s->peer_same_device[p];  // IPC peers: by the device tag behind their mailbox
            if (!s->peer_ipc[p]) {
                cudaPointerAttributes at;
                CU(ctx, cudaPointerGetAttributes(&at, s->peer[p]));
                same = at.device == ctx->device;
            }
            if (same)  // two ranks on one GPU: their barrier kernels could wait on each other forever
                return bsx_fail(ctx, BSX_ERR_UNSUPPORTED, "rank %d shares device %d with rank %d: the device-side barrier needs "
                                "one GPU per rank (use bsx_shard_transform / bsx_shard_commit with a host barrier)", p, ctx->device, s->rank);
        }
        M.mail[p] = shard_mail(s, p);
    }
    if (!s->timeout_flag) {
        CU(ctx, cudaHostAlloc((void **)&s->timeout_flag, sizeof(int), cudaHostAllocMapped));
        *s->timeout_flag = 0;
    }
    cudaEvent_t ev[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    if (timing)
        for (auto &e : ev) cudaEventCreate(&e);
    std::vector<cudaEvent_t> chunk_ev;
    const u64 epoch = ++s->epoch;
    int rc = shard_transform_enqueue(s, input, mem, input_is_coeffs, 0, timing ? ev : nullptr, nullptr, &chunk_ev);
    bsx_tree *t = nullptr;
    if (rc == BSX_OK) {
        BSX_COUNT_LAUNCH();
        shard_barrier_kernel<<<1, BSX_MAX_RANKS, 0, ctx->stream>>>(M, s->rank, s->n_ranks, epoch, 0, nullptr, 0, s->timeout_flag);
        if (timing) cudaEventRecord(ev[3], ctx->stream);
        t = new bsx_tree();
        t->ctx = ctx;
        t->log_leaves = s->log_n + s->rate_bits - s->log_ranks;
        t->leaf_len = s->n_cols_total;
        t->cap_height = s->cap_height > s->log_ranks ? s->cap_height - s->log_ranks : 0;
        t->leaves = s->slab;
        t->col_stride = s->chunk;
        t->owns_leaves = false;
        t->embedded = false;
        rc = bsx_tree_build(t, timing ? ev[4] : nullptr);
    }
    if (rc == BSX_OK) {
        BSX_COUNT_LAUNCH();
        shard_barrier_kernel<<<1, BSX_MAX_RANKS, 0, ctx->stream>>>(M, s->rank, s->n_ranks, epoch, 1, t->cap_dev, (int)s->cap_words,
                                                                  s->timeout_flag);
        if (timing) cudaEventRecord(ev[5], ctx->stream);
        rc = bsx_copy_out(ctx, (u64 *)rank_caps_host, M.mail[s->rank] + 2 * s->n_ranks, (size_t)s->n_ranks * s->cap_words, BSX_HOST);
    }
    cudaError_t e = cudaStreamSynchronize(ctx->stream);
    if (rc == BSX_OK && e != cudaSuccess) rc = bsx_fail(ctx, BSX_ERR_CUDA, "sharded commit failed: %s", cudaGetErrorString(e));
    if (rc == BSX_OK && *s->timeout_flag) {
        *s->timeout_flag = 0;
        rc = bsx_fail(ctx, BSX_ERR_CUDA, "a peer rank did not reach the device barrier within its time limit");
    }
    for (auto &x : chunk_ev) cudaEventDestroy(x);
    if (timing) {
        memset(timing, 0, sizeof(*timing));
        if (rc == BSX_OK && s->c1 > s->c0) {
            cudaEventElapsedTime(&timing->ifft_ms, ev[0], ev[1]);
            cudaEventElapsedTime(&timing->lde_ms, ev[1], ev[2]);
            cudaEventElapsedTime(&timing->h2d_ms, ev[2], ev[3]);  // here: the wait for the slowest peer's stores
            cudaEventElapsedTime(&timing->leaf_hash_ms, ev[3], ev[4]);
            cudaEventElapsedTime(&timing->tree_ms, ev[4], ev[5]);
            cudaEventElapsedTime(&timing->total_ms, ev[0], ev[5]);
        }
        for (auto &x : ev) cudaEventDestroy(x);
    }
    if (rc != BSX_OK) {
        if (t) {
            bsx_tree_release(t);
            delete t;
        }
        return rc;
    }
    *tree_out = t;
    return BSX_OK;
}

// Leaf hashing and the local cap subtrees over this rank's row slab (all slabs must be complete: the
// host puts a barrier, or the all-to-all, between bsx_shard_transform and this call).  The tree borrows
// the slab; local cap height = max(cap_height - log2(n_ranks), 0).
extern "C" int32_t bsx_shard_commit(bsx_shard *s, bsx_tree **out) {
    if (!s || !out) return BSX_ERR_INVALID_ARG;
    int local_cap = s->cap_height - s->log_ranks;
    if (local_cap < 0) local_cap = 0;
    return bsx_merkle_new_colmajor_device(s->ctx, (const uint64_t *)s->slab, s->chunk,
                                          s->log_n + s->rate_bits - s->log_ranks, s->n_cols_total, local_cap, out);
}

extern "C" int32_t bsx_shard_shape(const bsx_shard *s, int32_t *c0, int32_t *c1, uint64_t *chunk, int32_t *local_cap) {
    if (!s) return BSX_ERR_INVALID_ARG;
    if (c0) *c0 = s->c0;
    if (c1) *c1 = s->c1;
    if (chunk) *chunk = s->chunk;
    if (local_cap) *local_cap = s->cap_height > s->log_ranks ? s->cap_height - s->log_ranks : 0;
    return BSX_OK;
}

// polynomials[c0..c1) of this rank in natural order, column-major
extern "C" int32_t bsx_shard_coeffs(bsx_shard *s, uint64_t *out, int32_t mem) {
    if (!s) return BSX_ERR_INVALID_ARG;
    bsx_ctx *ctx = s->ctx;
    BsxScope scope(ctx);
    size_t n = (size_t)1 << s->log_n, words = n * (size_t)(s->c1 - s->c0);
    if (words == 0) return BSX_OK;
    if (!out || !s->coeffs) return BSX_ERR_INVALID_ARG;
    CU(ctx, cudaSetDevice(ctx->device));
    if (s->coeffs_bitrev) {
        u64 *tmp;
        TRY(bsx_alloc(ctx, &tmp, words));
        launch_bitrev_copy(s->coeffs, n, tmp, n, s->log_n, s->c1 - s->c0, ctx->stream);
        TRY(bsx_copy_out(ctx, (u64 *)out, tmp, words, mem));
        bsx_free(ctx, tmp);
    } else {
        TRY(bsx_copy_out(ctx, (u64 *)out, s->coeffs, words, mem));
    }
    CU(ctx, cudaStreamSynchronize(ctx->stream));
    return BSX_OK;
}

// n_in digests (power of two) hashed pairwise `levels` times: the top of the tree when there are more
// ranks than cap entries (cap_height < log2(n_ranks)), and a general two_to_one helper for hosts.
extern "C" int32_t bsx_merkle_reduce(bsx_ctx *ctx, const uint64_t *digests_host, int32_t n_in, int32_t levels,
                                     uint64_t *out_host) {
    if (!ctx || !digests_host || !out_host || n_in <= 0 || levels < 0 || (n_in >> levels) == 0 ||
        ((n_in >> levels) << levels) != n_in)
        return BSX_ERR_INVALID_ARG;
    BsxScope scope(ctx);
    CU(ctx, cudaSetDevice(ctx->device));
    u64 *a, *b;
    TRY(bsx_alloc(ctx, &a, (size_t)4 * n_in));
    TRY(bsx_alloc(ctx, &b, (size_t)4 * n_in));
    CU(ctx, cudaMemcpyAsync(a, digests_host, (size_t)32 * n_in, cudaMemcpyHostToDevice, ctx->stream));
    size_t cur = (size_t)n_in;
    for (int l = 0; l < levels; l++) {
        launch_merkle_level(a, b, cur >> 1, ctx->stream);
        u64 *t = a;
        a = b;
        b = t;
        cur >>= 1;
    }
    CU(ctx, cudaMemcpyAsync(out_host, a, 32 * cur, cudaMemcpyDeviceToHost, ctx->stream));
    bsx_free(ctx, a);
    bsx_free(ctx, b);
    CU(ctx, cudaStreamSynchronize(ctx->stream));
    return BSX_OK;
}

// prove_openings over row-sharded oracles, step 1 ("run FRI after gathering the small final-poly
// inputs").  `instance` is the instance of the whole proof: poly_oracle indexes `shards`, poly_index is the GLOBAL
// column of that oracle.  This rank contributes the polynomials whose coefficient columns it holds; the parts of all
// ranks add up (bsx_field_sum) to the final polynomial, which then goes to bsx_fri_proof_from_final_poly.
extern "C" int32_t bsx_shard_final_poly_part(bsx_shard *const *shards, int32_t n_shards, const bsx_fri_instance *instance,
                                             const uint64_t alpha[2], uint64_t *out, int32_t mem_out) {
    if (!shards || n_shards <= 0 || n_shards > 64 || !instance || !alpha || !out || !shards[0]) return BSX_ERR_INVALID_ARG;
    bsx_ctx *ctx = shards[0]->ctx;
    if (instance->n_batches < 0 || instance->n_batches > 16) return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "bad instance");
    std::vector<bsx_batch> views((size_t)n_shards);
    std::vector<bsx_batch *> ptrs((size_t)n_shards);
    for (int o = 0; o < n_shards; o++) {
        const bsx_shard *s = shards[o];
        if (!s || s->ctx != ctx || s->log_n != shards[0]->log_n)
            return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "shard %d does not match the first one", o);
        if (s->c1 > s->c0 && !s->coeffs) return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "shard %d has not been transformed", o);
        views[o].ctx = ctx;
        views[o].log_n = s->log_n;
        views[o].n_cols = s->c1 - s->c0;
        views[o].n_salt = 0;
        views[o].rate_bits = s->rate_bits;
        views[o].coeffs = s->coeffs;
        views[o].coeffs_bitrev = s->coeffs_bitrev;
        views[o].lde = nullptr;
        ptrs[o] = &views[o];
    }
    int total = 0;
    for (int b = 0; b < instance->n_batches; b++) {
        if (instance->n_polys[b] < 0) return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "bad instance");
        total += instance->n_polys[b];
    }
    std::vector<int32_t> po((size_t)total), pi((size_t)total);
    for (int i = 0; i < total; i++) {
        const int o = instance->poly_oracle[i], j = instance->poly_index[i];
        if (o < 0 || o >= n_shards || j < 0 || j >= shards[o]->n_cols_total)
            return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "instance names polynomial %d of oracle %d, which does not exist", j, o);
        const bool mine = j >= shards[o]->c0 && j < shards[o]->c1;
        po[i] = mine ? o : -1;
        pi[i] = mine ? j - shards[o]->c0 : 0;
    }
    bsx_fri_instance local = *instance;
    local.poly_oracle = po.data();
    local.poly_index = pi.data();
    return bsx_fri_final_poly_part(ctx, ptrs.data(), n_shards, &local, alpha, shards[0]->log_n, out, mem_out);
}
