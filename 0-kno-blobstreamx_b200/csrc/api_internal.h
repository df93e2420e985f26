// api_internal.h -- host-side structs behind the opaque handles of include/bsx_b200.h.
#pragma once
#include <map>
#include <utility>
#include <vector>

#include "../../include/bsx_b200.h"
#include "gl.cuh"
#include "internal.h"

#define BSX_STAGE_SLOTS 4

struct CosetKey {
    int log_n, r_log;
    u64 shift;
    bool operator<(const CosetKey &o) const {
        if (log_n != o.log_n) return log_n < o.log_n;
        if (r_log != o.r_log) return r_log < o.r_log;
        return shift < o.shift;
    }
};

struct bsx_ctx {
    int device;
    cudaStream_t stream;
    cudaStream_t copy_stream;  // host -> device copies of the pipelined from_values path
    char err[512];
    NttTables tables;
    std::map<int, PowTable> pow_tables;
    std::map<CosetKey, std::pair<const u64 *, const u64 *>> coset_tables;
    std::map<CosetKey, const u64 *> post_tables;  // key.shift = 0: inverse transform, no coset
    // scratch allocations of the ABI call in progress (see BsxScope): pointer + sequence number
    std::vector<std::pair<u64 *, unsigned long long>> scratch;
    unsigned long long scratch_seq = 0;
    int scope_depth = 0;
    // pinned staging ring for witnesses in pageable host memory (api_batch.cu HostStager); grown on demand
    u64 *stage_ring = nullptr;
    size_t stage_slot_words = 0;
    cudaEvent_t stage_done[BSX_STAGE_SLOTS] = {};
    // exact-size recycling of large device buffers (api.cu bsx_alloc_owned)
    std::map<u64 *, size_t> big_live;
    std::vector<std::pair<u64 *, size_t>> big_cache;
    size_t big_cache_bytes = 0;
};

// plonky2 hash/merkle_tree.rs MerkleTree { leaves, digests, cap }, device resident.
struct bsx_tree {
    bsx_ctx *ctx;
    int log_leaves, leaf_len, cap_height;
    const u64 *leaves;  // column-major: (row, col) at leaves[col * col_stride + row]
    size_t col_stride;
    bool owns_leaves;
    u64 *digests;  // level order, leaf digests first; 2 * (n_leaves - 2^cap_height) digests
    u64 *cap_dev;  // 2^cap_height digests
    std::vector<u64> cap_host;
    bool embedded;  // lives inside a bsx_batch
};

// plonky2 fri/oracle.rs PolynomialBatch { polynomials, merkle_tree, degree_log, rate_bits, blinding }
struct bsx_batch {
    bsx_ctx *ctx;
    int log_n, n_cols, n_salt, rate_bits;
    u64 *coeffs;        // n_cols x 2^log_n, column-major
    bool coeffs_bitrev;  // coefficient i stored at position bitrev(i)
    u64 *lde;           // (n_cols + n_salt) x 2^(log_n + rate_bits), column-major, leaf order
    bsx_tree tree;
};

int bsx_fail(bsx_ctx *ctx, int code, const char *fmt, ...);
// bsx_alloc: temporaries of the running ABI call -- whatever is still allocated when the call's BsxScope ends
// (an early return on an error path) is freed there.  bsx_alloc_owned: buffers that outlive the call (batch, tree,
// table and shard members); their owners free them.  bsx_free takes either kind.
int bsx_alloc(bsx_ctx *ctx, u64 **p, size_t words);
int bsx_alloc_owned(bsx_ctx *ctx, u64 **p, size_t words);
void bsx_free(bsx_ctx *ctx, u64 *p);
void bsx_ctx_drop_cache(bsx_ctx *ctx);
struct BsxScope {
    bsx_ctx *ctx;
    unsigned long long first;
    explicit BsxScope(bsx_ctx *c) : ctx(c), first(c ? c->scratch_seq : 0) {
        if (ctx) ctx->scope_depth++;
    }
    ~BsxScope();
    BsxScope(const BsxScope &) = delete;
    BsxScope &operator=(const BsxScope &) = delete;
};
int bsx_copy_in(bsx_ctx *ctx, u64 *dst_dev, const u64 *src, size_t words, int mem);
int bsx_copy_out(bsx_ctx *ctx, u64 *dst, const u64 *src_dev, size_t words, int mem);
int bsx_pow_table(bsx_ctx *ctx, int log_n, int inverse, PowTable *out);
int bsx_coset_tables(bsx_ctx *ctx, int log_n, int r_log, u64 shift, const u64 **pre_hi, const u64 **pre_lo);
// full pass-HI output-factor table for a plan whose pre_* / tw / scale / a / b / n_cosets are set; sets
// plan->post_full (left null when the table would not pay off: few columns, single pass, or too large)
int bsx_post_table(bsx_ctx *ctx, NttPlan *plan, u64 shift, int n_cols);
void bsx_split(int log_n, int *a, int *b);
void ntt_init_device();

struct BsxDests {
    int n, lc;  // n row slabs of 2^lc leaves each
    u64 *ptr[BSX_MAX_RANKS];
};
int bsx_commit_transforms(bsx_ctx *ctx, u64 *coeffs, int log_n, int n_cols, int rate_bits, bool input_is_coeffs,
                          u64 *lde, size_t lde_stride, const BsxDests *dests, cudaEvent_t ev_after_ifft);

int bsx_commit_ifft(bsx_ctx *ctx, u64 *coeffs, int log_n, int n_cols);
int bsx_commit_lde(bsx_ctx *ctx, const u64 *coeffs, bool coeffs_bitrev, int log_n, int n_cols, int rate_bits, u64 *lde,
                   size_t lde_stride, const BsxDests *dests);
int bsx_tree_alloc(bsx_tree *t);
void bsx_tree_hash_leaves(bsx_tree *t, size_t first, size_t count, cudaStream_t st);
int bsx_tree_finish(bsx_tree *t);

// builds digests + cap for a tree whose leaves/shape fields are set; allocates digests/cap_dev
int bsx_tree_build(bsx_tree *t, cudaEvent_t ev_after_leaf_hash);
void bsx_tree_release(bsx_tree *t);

#define CU(ctx, expr)                                                                             \
    do {                                                                                          \
        cudaError_t _e = (expr);                                                                  \
        if (_e != cudaSuccess) {                                                                  \
            cudaGetLastError();                                                                   \
            return bsx_fail((ctx), BSX_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), \
                            __FILE__, __LINE__);                                                  \
        }                                                                                         \
    } while (0)
#define TRY(expr)                  \
    do {                           \
        int _rc = (expr);          \
        if (_rc != BSX_OK) return _rc; \
    } while (0)
