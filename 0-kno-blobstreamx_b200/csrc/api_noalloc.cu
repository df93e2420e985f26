// api_noalloc.cu -- one entry point that allocates nothing: bsx_merkle_reduce_noalloc.
//
// No counterpart in the reference (plonky2 has no such call); added for callers
// that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
// It is bsx_merkle_reduce (api_shard.cu: n_in digests hashed pairwise `levels` times with PoseidonHash::two_to_one)
// with its two ping-pong scratch buffers taken from ONE file-scope device array instead of the stream-ordered pool.
//
//  * The array is BSX_NOALLOC_MAX_DIGESTS * 6 words = 3 MiB; every GPU that loads the library holds it until exit.
//  * In host code `bsx_noalloc_scratch` is a host-side stand-in, not a device address: the address is looked up once
//    per device with cudaGetSymbolAddress while that device is current, and only that pointer goes to the kernels.
//  * Every call on a device shares the array and finds the previous call's contents in it.  Nothing has to start at
//    zero: every scratch word a call reads was written earlier in the same call (the input copy or the level below),
//    so nothing is cleared.  Calls that overlap on one GPU (two contexts, two host threads) race on the array: the
//    caller serialises them.  The call is synchronous at return, like the rest of the ABI.
//  * No cudaMalloc / cudaFree / cudaMallocAsync and no allocating library routine on this path; larger inputs are
//    rejected with BSX_ERR_UNSUPPORTED, never served from the pool.
#include <mutex>

#include "api_internal.h"

#define NOALLOC_A_WORDS ((size_t)BSX_NOALLOC_MAX_DIGESTS * 4)  // inputs staged from the host / even levels
#define NOALLOC_B_WORDS ((size_t)BSX_NOALLOC_MAX_DIGESTS * 2)  // odd levels (at most half the input)
static __device__ __align__(256) u64 bsx_noalloc_scratch[NOALLOC_A_WORDS + NOALLOC_B_WORDS];

#define NOALLOC_MAX_DEVICES 64
static u64 *g_scratch_dev[NOALLOC_MAX_DEVICES];
static std::mutex g_scratch_mutex;

// device address of the array on ctx->device (that device must be current)
static int noalloc_scratch(bsx_ctx *ctx, u64 **out) {
    if (ctx->device < 0 || ctx->device >= NOALLOC_MAX_DEVICES)
        return bsx_fail(ctx, BSX_ERR_UNSUPPORTED, "device %d outside [0, %d)", ctx->device, NOALLOC_MAX_DEVICES);
    std::lock_guard<std::mutex> lk(g_scratch_mutex);
    if (!g_scratch_dev[ctx->device]) {
        void *p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, bsx_noalloc_scratch);
        if (e != cudaSuccess || !p) {
            cudaGetLastError();
            return bsx_fail(ctx, BSX_ERR_CUDA, "cudaGetSymbolAddress(bsx_noalloc_scratch) failed on device %d: %s", ctx->device,
                            cudaGetErrorString(e));
        }
        g_scratch_dev[ctx->device] = (u64 *)p;
    }
    *out = g_scratch_dev[ctx->device];
    return BSX_OK;
}

extern "C" int32_t bsx_merkle_reduce_noalloc(bsx_ctx *ctx, const uint64_t *digests, int32_t mem_in, int32_t n_in,
                                             int32_t levels, uint64_t *out, int32_t mem_out) {
    if (!ctx) return BSX_ERR_INVALID_ARG;
    if (!digests || !out || n_in <= 0 || levels < 0 || levels > 30 || (n_in >> levels) == 0 ||
        ((n_in >> levels) << levels) != n_in)
        return bsx_fail(ctx, BSX_ERR_INVALID_ARG, "bsx_merkle_reduce_noalloc: n_in=%d must be a positive multiple of 2^levels (levels=%d)",
                        n_in, levels);
    if (n_in > BSX_NOALLOC_MAX_DIGESTS)
        return bsx_fail(ctx, BSX_ERR_UNSUPPORTED, "bsx_merkle_reduce_noalloc: n_in=%d digests exceed the static scratch of "
                        "BSX_NOALLOC_MAX_DIGESTS=%d (use bsx_merkle_reduce, which allocates)", n_in, BSX_NOALLOC_MAX_DIGESTS);
    CU(ctx, cudaSetDevice(ctx->device));
    u64 *scratch = nullptr;
    TRY(noalloc_scratch(ctx, &scratch));
    u64 *A = scratch, *B = scratch + NOALLOC_A_WORDS;
    const u64 *cur = (const u64 *)digests;
    if (mem_in != BSX_DEVICE) {  // host digests: staged in A (an async copy from the caller's buffer allocates nothing)
        CU(ctx, cudaMemcpyAsync(A, digests, (size_t)32 * n_in, cudaMemcpyHostToDevice, ctx->stream));
        cur = A;
    }
    size_t n = (size_t)n_in;
    for (int l = 0; l < levels; l++) {
        u64 *dst = (cur == B) ? A : B;  // level 0 reads the caller's buffer or A and writes B
        launch_merkle_level(cur, dst, n >> 1, ctx->stream);
        cur = dst;
        n >>= 1;
    }
    CU(ctx, cudaGetLastError());
    CU(ctx, cudaMemcpyAsync(out, cur, 32 * n, mem_out == BSX_DEVICE ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost,
                            ctx->stream));
    CU(ctx, cudaStreamSynchronize(ctx->stream));
    return BSX_OK;
}
