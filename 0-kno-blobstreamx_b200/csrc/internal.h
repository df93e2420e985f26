// internal.h -- launcher prototypes shared by the translation units of libbsx_b200.so.
// Everything here is device-pointer based; the C ABI in include/bsx_b200.h sits on top (api.cu).
#pragma once
#include <cuda_runtime.h>
#include <cstddef>
#include <cstdint>

typedef unsigned long long u64;
typedef unsigned int u32;

// process-wide count of kernel launches issued by this library (bench.py reports it)
extern unsigned long long g_bsx_launches;
// relaxed atomic: several host threads (one per rank of a bsx_comm) launch concurrently
#define BSX_COUNT_LAUNCH() ((void)__atomic_fetch_add(&g_bsx_launches, 1ULL, __ATOMIC_RELAXED))

// ---------------------------------------------------------------- hashing / Merkle
// Leaves are COLUMN-major on device: element (row, col) at data[col * col_stride + row].
// digests_out: n_rows x 4 words.  Leaves of <= 4 elements are zero-extended (hash_or_noop).
void launch_leaf_hash(const u64 *data, size_t col_stride, int n_cols, size_t n_rows, u64 *digests_out,
                      cudaStream_t st);
// parents[i] = two_to_one(children[2i], children[2i+1]), n_parents digests
void launch_merkle_level(const u64 *children, u64 *parents, size_t n_parents, cudaStream_t st);
// Whole tree below the cap from leaf digests: level-order buffer (leaf level first), cap separately.
// digests holds 2*(n_leaves - 2^cap_height) digests with level 0 already filled when levels > 0;
// when log_leaves == cap_height the leaf digests ARE the cap and are expected in `cap`.
void launch_merkle_tree(u64 *digests, u64 *cap, int log_leaves, int cap_height, cudaStream_t st);
// states: n x 12 row-major, permuted in place (test hook)
void launch_poseidon_permute(u64 *states, size_t n, cudaStream_t st);
// row-major (n_rows x n_cols) <-> column-major (n_cols x n_rows, stride n_rows)
void launch_transpose(const u64 *src, u64 *dst, size_t n_rows, size_t n_cols, cudaStream_t st);
void launch_transpose_strided(const u64 *src, size_t src_stride, u64 *dst, size_t n_rows, size_t n_cols, cudaStream_t st);
// out[r * n_cols + c] = data[c * col_stride + idx[r]]
void launch_gather_rows(const u64 *data, size_t col_stride, int n_cols, const u64 *idx, int n_idx, u64 *out,
                        cudaStream_t st);
// siblings of leaf_index, bottom-up, from the level-order digest buffer: n_levels x 4 words
void launch_gather_siblings(const u64 *digests, int log_leaves, int n_levels, u64 leaf_index, u64 *out,
                            cudaStream_t st);

// ---------------------------------------------------------------- NTT
#define BSX_MAX_RANKS 16
struct NttTables {
    // omega_{4096}^j, j < 4096 (forward / inverse): sub-transform twiddles
    const u64 *w_fwd, *w_inv;
};
// base^e = hi[e >> 12] * lo[e & 4095], e < 2^24
struct PowTable {
    const u64 *lo, *hi;
};
struct NttPlan {
    int log_n;     // transform size
    int a, b;      // log_n = a + b: a high bits (strided pass), b low bits (contiguous pass)
    int inverse;   // use omega^-1 and scale by n^-1
    int src_bitrev;  // source holds coefficient i at position bitrev(i)
    int n_cosets;  // 1, or 2^rate_bits output cosets (LDE)
    // coset k pre-factor g_k^i, i = i1 * 2^b + i0: pre_hi[k << a | i1] * pre_lo[k << b | i0]; null = none.
    // pre_hi is followed by a second copy with every row bit-reversed: pre_hi[(n_cosets + k) << a | bitrev_a(i1)]
    const u64 *pre_hi, *pre_lo;
    PowTable tw;   // powers of omega_N (or its inverse) for the inter-pass twiddle
    u64 scale;     // multiplied into every output (n^-1 for the inverse); 1 = none
    // optional [n_cosets][2^a][2^b] table of scale * pre_lo * inter-pass twiddle (launch_ntt_post_table):
    // one multiplication per output word in pass HI instead of three, and no scaling in pass LO
    const u64 *post_full;
    // Row-sharded output (multi-GPU commit): when n_dest > 0 the word at leaf position l of column c is
    // stored to dest[l >> dest_lc] + c * 2^dest_lc + (l & (2^dest_lc - 1)) instead of dst + c * dst_stride
    // + l; the pointers may be peer-GPU memory (NVLink P2P stores) or slices of a local send buffer.
    // dst then only carries the intermediate between the two passes.
    int n_dest, dest_lc;
    u64 *dest[BSX_MAX_RANKS];
};
// Transforms n_cols columns.  src column c at src + c*src_stride (2^log_n words);
// dst column c at dst + c*dst_stride; coset k is written to the block [bitrev_r(k) << log_n, +2^log_n)
// of the column, every block in bit-reversed (DIF) order.  In place allowed when n_cosets == 1.
void launch_ntt(const NttTables &T, const NttPlan &plan, const u64 *src, size_t src_stride, u64 *dst,
                size_t dst_stride, int n_cols, cudaStream_t st);
// out[i] = first * base^(i * step_mul) for i < n
void launch_fill_powers(u64 *out, u64 base, u64 first, size_t n, cudaStream_t st);
void launch_scale_two_level(u64 *data, size_t stride, int log_n, int b, int n_cols, const u64 *hi, const u64 *lo,
                            cudaStream_t st);
void launch_any_nonzero(const u64 *data, size_t stride, size_t begin, size_t end, int n_cols, int *flag, cudaStream_t st);
// dst[c][i] = src[c][bitrev_log_n(i)]
void launch_bitrev_copy(const u64 *src, size_t src_stride, u64 *dst, size_t dst_stride, int log_n, int n_cols,
                        cudaStream_t st);
void launch_ntt_post_table(u64 *out, const u64 *pre_lo, PowTable tw, u64 scale, int a, int b, int r_log,
                           cudaStream_t st);
int ntt_max_single_pass_log();
void ntt_init_device();
void launch_fill_random(u64 *out, size_t n, const u32 key[8], int raw, cudaStream_t st);

// ---------------------------------------------------------------- FRI / challenger (fri_kernels.cu)
void launch_chal_observe(void *chal_dev, const u64 *data, size_t n, int mode, size_t plane_stride, cudaStream_t st);
void launch_chal_get(void *chal_dev, u64 *out, int n, u64 modulus, cudaStream_t st);
void launch_pow(const void *chal_dev, u64 base, u64 n, int pow_bits, u64 *best, cudaStream_t st);
void launch_chal_pow_finish(void *chal_dev, const u64 *witness, u64 *response, cudaStream_t st);
void launch_ext_powers(const u64 *alpha, u64 *pw, int n, cudaStream_t st);
void launch_combine(const u64 *const *ptrs, const int *pw_index, int n_polys, const u64 *pw, size_t n, u64 *comp0,
                    u64 *comp1, int accumulate, cudaStream_t st);
size_t divide_by_linear_workspace_words(size_t n);
void launch_divide_by_linear_accumulate(const u64 *comp0, const u64 *comp1, size_t n, const u64 *z, const u64 *sh,
                                        u64 *fin0, u64 *fin1, u64 *work, cudaStream_t st);
void launch_fri_fold(const u64 *in0, const u64 *in1, size_t n_out, int arity_bits, const u64 *beta, u64 *out0,
                     u64 *out1, cudaStream_t st);
void launch_fri_leaves(const u64 *v0, const u64 *v1, size_t len, int arity_bits, u64 *leaves, cudaStream_t st);
void launch_fri_query_gather(const u64 *idx, int n_queries, int shift, const u64 *leaves, size_t col_stride, int width,
                             const u64 *digests, int log_leaves, int n_sib, u64 *out, size_t q_stride, size_t off,
                             cudaStream_t st);
void launch_ext_interleave(const u64 *planar, size_t plane_stride, size_t n, u64 *out, cudaStream_t st);
void launch_ext_deinterleave(const u64 *in, size_t n, u64 *planar, size_t plane_stride, cudaStream_t st);
void launch_ext_pow(const u64 *a, u64 k, u64 *out, cudaStream_t st);
void launch_field_sum(const u64 *parts, int k, size_t words, u64 *out, cudaStream_t st);

// ---------------------------------------------------------------- opening-set evaluation (eval_kernels.cu)
#define BSX_EVAL_MAX_BITS 64
size_t eval_workspace_words(int log_n, int n_cols, int n_points);
void launch_eval_ext(const u64 *coeffs, size_t col_stride, int log_n, bool bitrev, int n_cols, const u64 *points,
                     int n_points, u64 *work, u64 *out, cudaStream_t st);
