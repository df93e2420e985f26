// hash_kernels.cu -- Poseidon leaf hashing and Merkle tree kernels (K4, K5, K8 of SURVEY.md).
// Replaces plonky2 hash/merkle_tree.rs (`MerkleTree::new`, `fill_digests_buf`, `fill_subtree`,
// `prove`, `get`) and hash/hashing.rs (`hash_or_noop`, `hash_n_to_hash_no_pad`).
#include "internal.h"
#include "poseidon.cuh"

// 64 registers, four blocks of 256 per SM.  The kernel is bound by integer issue, eight warps per scheduler
// hide the column loads; a register prefetch of the next chunk cost occupancy (104.0 ms against 102.2 ms).
#ifndef LEAF_BLOCKS
#define LEAF_BLOCKS 4
#endif

// ------------------------------------------------------------------ leaf hash
// One thread per leaf row.  The LDE buffer is column-major, so lane l of a warp reads word
// (col, row0 + l): 32 lanes x 8 B = 256 contiguous bytes per column -- the "transpose" of
// plonky2's fri/oracle.rs (transpose + reverse_index_bits_in_place) never materialises.
__global__ void __launch_bounds__(256, LEAF_BLOCKS)
    leaf_hash_kernel(const u64 *__restrict__ data, size_t col_stride, int n_cols, size_t n_rows, u64 *__restrict__ digests) {
    size_t row = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= n_rows) return;
    const u64 *p = data + row;
    u64 s[12];
#pragma unroll
    for (int i = 0; i < 12; i++) s[i] = 0;
    if (n_cols <= 4) {  // hash_or_noop: zero-extend, no permutation
#pragma unroll
        for (int c = 0; c < 4; c++)  // static indices: a run-time index would put the state in local memory
            if (c < n_cols) s[c] = p[(size_t)c * col_stride];
    } else {
        // One permutation call site (code size).  A short last chunk overwrites only its own lanes (no padding,
        // hashing.rs): its missing columns re-read the last one and are dropped by a select, so the loads stay
        // unconditional and the state never leaves the registers (predicated loads parked it on the stack).
        const int last = n_cols - 1;
#pragma unroll 1
        for (int c = 0; c < n_cols; c += 8) {
            u64 v[8];
#pragma unroll
            for (int i = 0; i < 8; i++) v[i] = __ldg(p + (size_t)min(c + i, last) * col_stride);
#pragma unroll
            for (int i = 0; i < 8; i++) s[i] = (c + i <= last) ? v[i] : s[i];
            poseidon_permute_lazy(s);
        }
#pragma unroll
        for (int i = 0; i < 4; i++) s[i] = gl_canon(s[i]);
    }
    ulonglong2 *o = reinterpret_cast<ulonglong2 *>(digests + 4 * row);
    o[0] = make_ulonglong2(s[0], s[1]);
    o[1] = make_ulonglong2(s[2], s[3]);
}

void launch_leaf_hash(const u64 *data, size_t col_stride, int n_cols, size_t n_rows, u64 *digests_out,
                      cudaStream_t st) {
    if (n_rows == 0) return;
    unsigned blocks = (unsigned)((n_rows + 255) / 256);
    BSX_COUNT_LAUNCH(); leaf_hash_kernel<<<blocks, 256, 0, st>>>(data, col_stride, n_cols, n_rows, digests_out);
}

// ------------------------------------------------------------------ inner levels
__global__ void __launch_bounds__(256) merkle_level_kernel(const u64 *__restrict__ children, u64 *__restrict__ parents,
                                                           size_t n_parents) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_parents) return;
    const ulonglong2 *c = reinterpret_cast<const ulonglong2 *>(children + 8 * i);
    ulonglong2 a = c[0], b = c[1], d = c[2], e = c[3];
    u64 s[12] = {a.x, a.y, b.x, b.y, d.x, d.y, e.x, e.y, 0, 0, 0, 0};
    poseidon_permute_lazy(s);
#pragma unroll
    for (int t = 0; t < 4; t++) s[t] = gl_canon(s[t]);
    ulonglong2 *o = reinterpret_cast<ulonglong2 *>(parents + 4 * i);
    o[0] = make_ulonglong2(s[0], s[1]);
    o[1] = make_ulonglong2(s[2], s[3]);
}

void launch_merkle_level(const u64 *children, u64 *parents, size_t n_parents, cudaStream_t st) {
    if (n_parents == 0) return;
    unsigned blocks = (unsigned)((n_parents + 255) / 256);
    BSX_COUNT_LAUNCH(); merkle_level_kernel<<<blocks, 256, 0, st>>>(children, parents, n_parents);
}

// One block per subtree of 2^SUB_LOG nodes at the input level: digest layers stay in shared
// memory while the block climbs `levels` levels; every layer is also written to its level-order
// slot in global memory because `MerkleTree::prove` reads them later.
// (Round 2 tried the warp-wide permutation of poseidon.cuh for the narrow layers -- one parent per warp instead of per
// thread -- and a single launch up to the cap: tree 0.374 -> 0.439 ms at 2^17 leaves, 1.088 -> 1.153 ms at 2^20.  The
// naive shuffle form is slower per layer than one thread per parent on an otherwise idle SM; reverted.)
#define SUB_LOG 9
#define MERKLE_FLAT_MIN_LOG 15  // levels with more than 2^15 inputs run flat, the rest in per-block subtrees
__global__ void __launch_bounds__(256) merkle_subtree_kernel(const u64 *__restrict__ in_level, u64 *__restrict__ out0,
                                                             size_t n_in, int levels, size_t level_stride_words,
                                                             u64 *__restrict__ cap, int cap_level) {
    // in_level: n_in digests.  Block b owns inputs [b << SUB_LOG, (b+1) << SUB_LOG).
    // Level j (1-based) output goes to out_j = out0 + sum_{t<j-1} (n_in >> (t+1)) * 4 words, unless
    // j == cap_level, which goes to `cap`.
    __shared__ u64 sm[2][(1 << (SUB_LOG - 1)) * 4];
    size_t base_in = (size_t)blockIdx.x << SUB_LOG;
    size_t n_par = (size_t)1 << (SUB_LOG - 1);
    const u64 *src = in_level + 4 * base_in;
    u64 *out = out0;
    size_t lvl_n = n_in >> 1;  // digests in the level being produced (whole tree)
    int buf = 0;
    (void)level_stride_words;
    for (int j = 1; j <= levels; j++) {
        u64 *dst_g = (j == cap_level) ? cap : out;
        size_t base_out = ((size_t)blockIdx.x << SUB_LOG) >> j;
        for (size_t i = threadIdx.x; i < n_par; i += blockDim.x) {
            u64 s[12];
            if (j == 1) {
                const ulonglong2 *c = reinterpret_cast<const ulonglong2 *>(src + 8 * i);
                ulonglong2 a = c[0], b = c[1], d = c[2], e = c[3];
                s[0] = a.x; s[1] = a.y; s[2] = b.x; s[3] = b.y;
                s[4] = d.x; s[5] = d.y; s[6] = e.x; s[7] = e.y;
            } else {
                const u64 *c = sm[buf ^ 1] + 8 * i;
#pragma unroll
                for (int t = 0; t < 8; t++) s[t] = c[t];
            }
            s[8] = s[9] = s[10] = s[11] = 0;
            poseidon_permute_lazy(s);
#pragma unroll
            for (int t = 0; t < 4; t++) s[t] = gl_canon(s[t]);
            u64 *d_s = sm[buf] + 4 * i;
            d_s[0] = s[0]; d_s[1] = s[1]; d_s[2] = s[2]; d_s[3] = s[3];
            ulonglong2 *o = reinterpret_cast<ulonglong2 *>(dst_g + 4 * (base_out + i));
            o[0] = make_ulonglong2(s[0], s[1]);
            o[1] = make_ulonglong2(s[2], s[3]);
        }
        __syncthreads();
        out += 4 * lvl_n;
        lvl_n >>= 1;
        n_par >>= 1;
        buf ^= 1;
    }
}

void launch_merkle_tree(u64 *digests, u64 *cap, int log_leaves, int cap_height, cudaStream_t st) {
    int levels = log_leaves - cap_height;  // levels to produce above the leaf digests
    if (levels <= 0) return;
    size_t n = (size_t)1 << log_leaves;
    const u64 *in = digests;
    u64 *out = digests + 4 * n;
    int done = 0;
    size_t n_in = n;
    while (done < levels) {
        int rem = levels - done;
        int log_in = log_leaves - done;
        if (n_in > ((size_t)1 << MERKLE_FLAT_MIN_LOG) && rem >= 1) {
            // wide levels: one thread per parent keeps every lane busy (the subtree kernel idles 7 of 8
            // warps above its third level, which cost 8 ms at 2^23 leaves)
            size_t n_par = n_in >> 1;
            u64 *dst = (rem == 1) ? cap : out;
            launch_merkle_level(in, dst, n_par, st);
            in = dst;
            out = dst + 4 * n_par;
            n_in = n_par;
            done += 1;
        } else if (log_in >= SUB_LOG && rem >= 1) {
            int climb = rem < SUB_LOG ? rem : SUB_LOG;
            unsigned blocks = (unsigned)(n_in >> SUB_LOG);
            BSX_COUNT_LAUNCH(); merkle_subtree_kernel<<<blocks, 256, 0, st>>>(in, out, n_in, climb, 0, cap, rem <= SUB_LOG ? rem : -1);
            // advance: the last produced level becomes the next input
            size_t words = 0, ln = n_in >> 1;
            for (int j = 1; j < climb; j++) {
                words += 4 * ln;
                ln >>= 1;
            }
            u64 *last = out + words;  // level `climb` output (unless it went to cap, then loop ends)
            in = last;
            out = last + 4 * ln;
            n_in >>= climb;
            done += climb;
        } else {
            // small remainder: one launch per level
            size_t n_par = n_in >> 1;
            u64 *dst = (rem == 1) ? cap : out;
            launch_merkle_level(in, dst, n_par, st);
            in = dst;
            out = dst + 4 * n_par;
            n_in = n_par;
            done += 1;
        }
    }
}

// ------------------------------------------------------------------ test hook / accessors
__global__ void poseidon_permute_kernel(u64 *states, size_t n) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    u64 s[12];
#pragma unroll
    for (int t = 0; t < 12; t++) s[t] = gl_canon(states[12 * i + t]);
    poseidon_permute(s);
#pragma unroll
    for (int t = 0; t < 12; t++) states[12 * i + t] = s[t];
}
void launch_poseidon_permute(u64 *states, size_t n, cudaStream_t st) {
    if (n == 0) return;
    BSX_COUNT_LAUNCH(); poseidon_permute_kernel<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(states, n);
}

// 32x32 tiled transpose through shared memory, both sides coalesced.
__global__ void transpose_kernel(const u64 *__restrict__ src, size_t src_stride, u64 *__restrict__ dst, size_t n_rows,
                                 size_t n_cols) {
    __shared__ u64 tile[32][33];
    // tiles on a one-dimensional grid: either side may exceed the 65535 limit of grid.y
    const size_t tiles_x = (n_cols + 31) / 32;
    size_t c0 = (blockIdx.x % tiles_x) * 32, r0 = (blockIdx.x / tiles_x) * 32;
    for (int j = threadIdx.y; j < 32; j += blockDim.y) {
        size_t r = r0 + j, c = c0 + threadIdx.x;
        if (r < n_rows && c < n_cols) tile[j][threadIdx.x] = src[r * src_stride + c];
    }
    __syncthreads();
    for (int j = threadIdx.y; j < 32; j += blockDim.y) {
        size_t c = c0 + j, r = r0 + threadIdx.x;
        if (r < n_rows && c < n_cols) dst[c * n_rows + r] = tile[threadIdx.x][j];
    }
}
void launch_transpose(const u64 *src, u64 *dst, size_t n_rows, size_t n_cols, cudaStream_t st) {
    launch_transpose_strided(src, n_cols, dst, n_rows, n_cols, st);
}
// src row r at src + r * src_stride (n_cols words used) -> dst (n_cols x n_rows), dense
void launch_transpose_strided(const u64 *src, size_t src_stride, u64 *dst, size_t n_rows, size_t n_cols, cudaStream_t st) {
    if (n_rows == 0 || n_cols == 0) return;
    const size_t tiles = ((n_cols + 31) / 32) * ((n_rows + 31) / 32);
    BSX_COUNT_LAUNCH(); transpose_kernel<<<(unsigned)tiles, dim3(32, 8), 0, st>>>(src, src_stride, dst, n_rows, n_cols);
}

__global__ void gather_rows_kernel(const u64 *__restrict__ data, size_t col_stride, int n_cols,
                                   const u64 *__restrict__ idx, int n_idx, u64 *__restrict__ out) {
    int r = blockIdx.x;
    if (r >= n_idx) return;
    u64 row = idx[r];
    for (int c = threadIdx.x; c < n_cols; c += blockDim.x) out[(size_t)r * n_cols + c] = data[(size_t)c * col_stride + row];
}
void launch_gather_rows(const u64 *data, size_t col_stride, int n_cols, const u64 *idx, int n_idx, u64 *out,
                        cudaStream_t st) {
    if (n_idx == 0 || n_cols == 0) return;
    BSX_COUNT_LAUNCH(); gather_rows_kernel<<<n_idx, 128, 0, st>>>(data, col_stride, n_cols, idx, n_idx, out);
}

__global__ void gather_siblings_kernel(const u64 *__restrict__ digests, int log_leaves, int n_levels, u64 leaf_index,
                                       u64 *__restrict__ out) {
    int k = blockIdx.x, t = threadIdx.x;  // 4 threads per level
    if (k >= n_levels || t >= 4) return;
    size_t off = 0;
    for (int j = 0; j < k; j++) off += (size_t)1 << (log_leaves - j);
    size_t idx = (leaf_index >> k) ^ 1;
    out[4 * k + t] = digests[4 * (off + idx) + t];
}
void launch_gather_siblings(const u64 *digests, int log_leaves, int n_levels, u64 leaf_index, u64 *out,
                            cudaStream_t st) {
    if (n_levels <= 0) return;
    BSX_COUNT_LAUNCH(); gather_siblings_kernel<<<n_levels, 4, 0, st>>>(digests, log_leaves, n_levels, leaf_index, out);
}
