/* bsx_b200.h -- C ABI of libbsx_b200.so: the B200-native plonky2 commitment core.
 *
 * This is the drop-in boundary for the hot path of the BlobstreamX prover.
 * The reference has NO FFI for this path: plonky2 exposes plain Rust functions that its prover
 * calls directly.  Each entry point below names the upstream plonky2 function it replaces; the
 * reference mount (/root/reference) is empty, so upstream paths are cited instead of file:line
 * (SURVEY.md section 0).  INTEGRATION.md shows the Rust `extern "C"` block and the patched
 * `PolynomialBatch` / `MerkleTree` / `fri_proof` shims a maintainer would add.
 *
 * Conventions
 *  - every field element is a canonical little-endian u64 (< p = 2^64 - 2^32 + 1); extension
 *    elements are two consecutive u64 [c0, c1]; digests are 4 x u64.
 *  - every function returns 0 on success or a negative BSX_ERR_* code; bsx_last_error(ctx) gives
 *    the message.  Upstream panics on the same misuse (non power-of-two lengths,
 *    cap_height > log2(leaves)); the Rust shim turns a non-zero code into panic!().
 *  - one bsx_ctx per (host thread, device); calls on a ctx are serialised by the caller; distinct
 *    contexts are independent.  Functions are synchronous at return.
 *  - `mem` says where a caller buffer lives: BSX_HOST (pageable or pinned) or BSX_DEVICE.
 *  - there is no CPU fallback: without a CUDA device bsx_ctx_create fails.
 */
#ifndef BSX_B200_H
#define BSX_B200_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define BSX_OK 0
#define BSX_ERR_INVALID_ARG (-1)
#define BSX_ERR_CUDA (-2)
#define BSX_ERR_OOM (-3)
#define BSX_ERR_UNSUPPORTED (-4)

#define BSX_HOST 0
#define BSX_DEVICE 1

#define BSX_SALT_SIZE 4 /* plonky2 plonk/plonk_common.rs SALT_SIZE */
#define BSX_MAX_ORACLES 16 /* oracles per prove_openings call */

typedef struct bsx_ctx bsx_ctx;
typedef struct bsx_tree bsx_tree;   /* plonky2 hash/merkle_tree.rs MerkleTree<F, PoseidonHash> */
typedef struct bsx_batch bsx_batch; /* plonky2 fri/oracle.rs PolynomialBatch<F, C, 2> */

/* per-phase device milliseconds (CUDA events), the analogue of plonky2 util/timing.rs TimingTree
 * scopes "IFFT", "FFT + blinding", "build Merkle tree" inside from_values. */
typedef struct bsx_timing {
    float h2d_ms;
    float ifft_ms;
    float lde_ms;
    float leaf_hash_ms;
    float tree_ms;
    float total_ms;
} bsx_timing;

const char *bsx_version(void);
int32_t bsx_device_count(void);
/* kernels launched by this library in this process so far (bench.py `gpu_launches`) */
uint64_t bsx_launch_count(void);

int32_t bsx_ctx_create(int32_t device, bsx_ctx **out);
void bsx_ctx_destroy(bsx_ctx *ctx);
const char *bsx_last_error(const bsx_ctx *ctx);
/* the CUDA stream all work of this ctx runs on (cudaStream_t as void*), for event timing */
void *bsx_ctx_stream(const bsx_ctx *ctx);
int32_t bsx_ctx_sync(bsx_ctx *ctx);
/* Device memory currently handed out by the stream-ordered pool the library allocates from (bytes; batches,
 * trees, cached tables).  A failed call leaves this unchanged: its temporaries are released on every path. */
int32_t bsx_ctx_mem_used(bsx_ctx *ctx, uint64_t *bytes);
/* Large buffers (>= 64 MiB) released by batches and calls are kept by the context and reused by exact size, so that a
 * prover committing the same shapes proof after proof never waits for the allocator; bsx_ctx_trim gives them back to
 * the device (they are also given back automatically when an allocation would otherwise fail). */
int32_t bsx_ctx_trim(bsx_ctx *ctx);

/* ---- PolynomialBatch (fri/oracle.rs) ------------------------------------------------------
 * from_values: values = n_cols column vectors of 2^log_n evaluations on the subgroup, column-major
 * (column c at values + c * 2^log_n), i.e. Vec<PolynomialValues<F>>.  Does ifft per column,
 * zero-pad by 2^rate_bits, coset FFT with shift g = 7, appends BSX_SALT_SIZE salt columns iff
 * blinding, and builds the Poseidon Merkle tree over the bit-reversed LDE rows down to the cap.
 * salt: optional caller-supplied salt columns (BSX_SALT_SIZE x 2^(log_n+rate_bits), column-major,
 * LDE-domain natural order, same `mem`); NULL with blinding != 0 draws them on the device
 * (upstream uses F::rand_vec, so salted caps are not reproducible against any reference).
 * Replaces PolynomialBatch::from_values. */
int32_t bsx_batch_from_values(bsx_ctx *ctx, const uint64_t *values, int32_t mem, int32_t log_n, int32_t n_cols,
                              int32_t rate_bits, int32_t cap_height, int32_t blinding, const uint64_t *salt,
                              bsx_batch **out, bsx_timing *timing);
/* The same with one HOST pointer per column -- `Vec<PolynomialValues<F>>` as it lies in the caller's heap, every
 * column its own allocation, no gather on the caller's side.  Pageable memory is fine: host threads
 * (BSX_COPY_THREADS, default 8) stage the columns through a pinned ring owned by the context while the DMA engine
 * and the transforms of the finished column chunks run underneath. */
int32_t bsx_batch_from_values_cols(bsx_ctx *ctx, const uint64_t *const *cols_host, int32_t log_n, int32_t n_cols,
                                   int32_t rate_bits, int32_t cap_height, int32_t blinding, const uint64_t *salt_host,
                                   bsx_batch **out, bsx_timing *timing);
int32_t bsx_batch_from_coeffs_cols(bsx_ctx *ctx, const uint64_t *const *cols_host, int32_t log_n, int32_t n_cols,
                                   int32_t rate_bits, int32_t cap_height, int32_t blinding, const uint64_t *salt_host,
                                   bsx_batch **out, bsx_timing *timing);
/* Replaces PolynomialBatch::from_coeffs: coeffs = n_cols coefficient vectors of length 2^log_n. */
int32_t bsx_batch_from_coeffs(bsx_ctx *ctx, const uint64_t *coeffs, int32_t mem, int32_t log_n, int32_t n_cols,
                              int32_t rate_bits, int32_t cap_height, int32_t blinding, const uint64_t *salt,
                              bsx_batch **out, bsx_timing *timing);
void bsx_batch_free(bsx_batch *b);
/* merkle_tree.cap: 2^cap_height digests -> out_host[2^cap_height * 4] */
int32_t bsx_batch_cap(const bsx_batch *b, uint64_t *out_host);
/* the batch's MerkleTree (borrowed; freed with the batch) */
bsx_tree *bsx_batch_tree(bsx_batch *b);
/* Replaces PolynomialBatch::get_lde_values(index, step): row bitrev(index * step) of the leaves with
 * the salt stripped -> out[n_cols].  `index` counts in the degree * 2^log2(step)... domain exactly as
 * upstream: leaf = reverse_bits(index * step, log_n + rate_bits). */
int32_t bsx_batch_get_lde_values(bsx_batch *b, uint64_t index, uint64_t step, uint64_t *out_host);
/* Replaces PolynomialBatch::get_lde_values_packed(index_start, step): the `count` rows
 * get_lde_values(index_start + i, step), i < count, row-major -> out[count][n_cols] on the `mem` side
 * (upstream packs P::WIDTH of them into SIMD lanes; the quotient loop reads every row this way). */
int32_t bsx_batch_get_lde_values_packed(bsx_batch *b, uint64_t index_start, uint64_t count, uint64_t step,
                                        uint64_t *out, int32_t mem);
/* polynomials[c].coeffs in natural order -> out[n_cols * 2^log_n] (column-major), `mem` side */
int32_t bsx_batch_coeffs(bsx_batch *b, uint64_t *out, int32_t mem);
/* Full export for serialisation (circuit artefacts): any pointer may be NULL.
 * leaves: row-major 2^(log_n+rate_bits) x (n_cols+salt), rows in leaf (bit-reversed) order;
 * digests: plonky2's `digests` Vec layout (fill_subtree order), 2*(leaves - 2^cap_height) x 4 words. */
int32_t bsx_batch_export(bsx_batch *b, uint64_t *coeffs_host, uint64_t *leaves_host, uint64_t *digests_host);
/* The inverse of bsx_batch_export (upstream: the circuit-artefact reader restoring the constants + sigmas
 * PolynomialBatch, util/serialization `read_polynomial_batch`): same layouts as export, nothing is recomputed
 * or verified.  n_salt = 0 or 4 (`blinding`).  All four parts are required. */
int32_t bsx_batch_import(bsx_ctx *ctx, const uint64_t *coeffs_host, const uint64_t *leaves_host,
                         const uint64_t *digests_host, const uint64_t *cap_host, int32_t log_n, int32_t n_cols,
                         int32_t n_salt, int32_t rate_bits, int32_t cap_height, bsx_batch **out);
int32_t bsx_batch_shape(const bsx_batch *b, int32_t *log_n, int32_t *n_cols, int32_t *rate_bits,
                        int32_t *cap_height, int32_t *n_salt);
/* device pointers for callers that keep working on the GPU (quotient evaluation, multi-GPU exchange):
 * LDE is column-major, column c at lde + c * 2^(log_n+rate_bits), position l = leaf index. */
const uint64_t *bsx_batch_lde_device(const bsx_batch *b);
const uint64_t *bsx_batch_digests_device(const bsx_batch *b);

/* ---- MerkleTree (hash/merkle_tree.rs) ------------------------------------------------------
 * Replaces MerkleTree::new(leaves, cap_height): leaves row-major 2^log_leaves x leaf_len. */
int32_t bsx_merkle_new(bsx_ctx *ctx, const uint64_t *leaves, int32_t mem, int32_t log_leaves, int32_t leaf_len,
                       int32_t cap_height, bsx_tree **out);
/* Same tree from column-major device leaves (element (row, col) at leaves[col * col_stride + row]);
 * the tree BORROWS the leaves (no copy).  Used by the multi-GPU path after the all-to-all. */
int32_t bsx_merkle_new_colmajor_device(bsx_ctx *ctx, const uint64_t *leaves_dev, size_t col_stride,
                                       int32_t log_leaves, int32_t leaf_len, int32_t cap_height, bsx_tree **out);
void bsx_merkle_free(bsx_tree *t);
int32_t bsx_merkle_cap(const bsx_tree *t, uint64_t *out_host);
/* MerkleTree::get(i) for k indices: out_host[k * leaf_len] */
int32_t bsx_merkle_get(bsx_tree *t, const uint64_t *idx_host, int32_t k, uint64_t *out_host);
/* MerkleTree::prove(leaf_index).siblings bottom-up: out_host[(log_leaves - cap_height) * 4] */
int32_t bsx_merkle_prove(bsx_tree *t, uint64_t leaf_index, uint64_t *siblings_host);
/* digests in plonky2's Vec layout (host) */
int32_t bsx_merkle_digests(bsx_tree *t, uint64_t *out_host);
int32_t bsx_merkle_shape(const bsx_tree *t, int32_t *log_leaves, int32_t *leaf_len, int32_t *cap_height);

/* pairwise two_to_one reduction of n_in digests (host) `levels` times -> n_in >> levels digests (host).
 * The top of a tree assembled from per-rank roots; PoseidonHash::two_to_one for host code. */
int32_t bsx_merkle_reduce(bsx_ctx *ctx, const uint64_t *digests_host, int32_t n_in, int32_t levels,
                          uint64_t *out_host);

/* The same reduction WITHOUT any allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).
 * The ping-pong scratch is a static device array of the library (3 MiB per GPU that loads it, held until exit), so
 * n_in is limited to BSX_NOALLOC_MAX_DIGESTS; more fails with BSX_ERR_UNSUPPORTED and a message.  `digests` and `out`
 * may each be host or device memory of ctx's device (mem_in / mem_out).  All calls on one GPU share the scratch, which
 * keeps the previous call's contents (nothing in it needs clearing: every word read was written by the same call);
 * calls that overlap on one GPU race and must be serialised by the caller.  Synchronous at return. */
#define BSX_NOALLOC_MAX_DIGESTS 65536
int32_t bsx_merkle_reduce_noalloc(bsx_ctx *ctx, const uint64_t *digests, int32_t mem_in, int32_t n_in, int32_t levels,
                                  uint64_t *out, int32_t mem_out);

/* ---- one commit sharded over the GPUs of a box (no upstream counterpart; BASELINE.json north star) ------
 * PolynomialBatch::from_values split as: columns -> ranks for the transforms, leaf rows -> ranks for
 * hashing.  Rank g of G (a power of two, <= 16) owns columns bsx_shard_column_range() and the leaf rows
 * [g * chunk, (g+1) * chunk), chunk = 2^(log_n + rate_bits) / G, of every column, held in its "row slab"
 * (n_cols_total x chunk, column-major).  Call order on every rank:
 *   bsx_shard_create -> exchange slab addresses (bsx_shard_slab_ipc / bsx_shard_set_peer_ipc across
 *   processes, bsx_shard_set_peer_ptr inside one) -> bsx_shard_transform(use_staging = 0): the last NTT pass
 *   stores every LDE word into its owner's slab over NVLink -> host barrier -> bsx_shard_commit.
 * Without peer access: bsx_shard_transform(use_staging = 1) fills a destination-major staging buffer, the
 * host runs one all-to-all (block d of n_local_cols * chunk words goes to rank d's slab at offset
 * c0 * chunk words), then bsx_shard_commit.  The local caps, concatenated in rank order, are the cap of the
 * whole batch (followed by bsx_merkle_reduce when cap_height < log2 G); every word equals the
 * single-GPU result. */
typedef struct bsx_shard bsx_shard;
#define BSX_MAX_RANKS 16
#define BSX_IPC_HANDLE_BYTES 64 /* sizeof(cudaIpcMemHandle_t) */
int32_t bsx_shard_column_range(int32_t n_cols_total, int32_t rank, int32_t n_ranks, int32_t *c0, int32_t *c1);
int32_t bsx_shard_create(bsx_ctx *ctx, int32_t log_n, int32_t n_cols_total, int32_t rate_bits, int32_t cap_height,
                         int32_t rank, int32_t n_ranks, bsx_shard **out);
void bsx_shard_free(bsx_shard *s);
int32_t bsx_shard_shape(const bsx_shard *s, int32_t *c0, int32_t *c1, uint64_t *chunk, int32_t *local_cap_height);
uint64_t *bsx_shard_slab_device(bsx_shard *s);
uint64_t *bsx_shard_staging_device(bsx_shard *s);
const uint64_t *bsx_shard_coeffs_device(const bsx_shard *s);
int32_t bsx_shard_slab_ipc(bsx_shard *s, uint8_t *handle_out /* BSX_IPC_HANDLE_BYTES */);
int32_t bsx_shard_set_peer_ipc(bsx_shard *s, int32_t peer_rank, const uint8_t *handle);
int32_t bsx_shard_set_peer_ptr(bsx_shard *s, int32_t peer_rank, uint64_t *slab_dev);
/* input: this rank's columns only, (c1 - c0) x 2^log_n column-major values (or coefficients) */
int32_t bsx_shard_transform(bsx_shard *s, const uint64_t *input, int32_t mem, int32_t input_is_coeffs,
                            int32_t use_staging, bsx_timing *timing);
/* The transforms of the local columns [col_begin, col_begin + col_count) only (`input` = those columns): a host running
 * the staged exchange sends the finished columns of one chunk while the next chunk is transformed.  Call with increasing
 * col_begin starting at 0.  sync == 0 returns with the work enqueued on the context's stream. */
int32_t bsx_shard_transform_cols(bsx_shard *s, const uint64_t *input, int32_t mem, int32_t input_is_coeffs,
                                 int32_t use_staging, int32_t col_begin, int32_t col_count, int32_t sync);
/* The whole commit of one rank with no host synchronisation between its phases, for one process per GPU with every
 * peer slab mapped (bsx_shard_set_peer_ipc): transforms -> device-side barrier over flags in the peers' slabs -> leaf
 * hashing and cap subtrees -> the local caps published to every rank -> rank_caps_host (n_ranks x 2^(local cap height)
 * x 4 words, rank order; equal on every rank).  A pinned host input is uploaded in column chunks under the transforms.
 * Collective: every rank must call it the same number of times.  BSX_ERR_UNSUPPORTED when two ranks share a device
 * (their barrier kernels could wait on each other; detected for peer pointers and, through a device tag behind the slab,
 * for IPC peers of other processes): use bsx_shard_transform / host barrier / bsx_shard_commit there.
 * timing->h2d_ms carries the wait for the slowest peer. */
int32_t bsx_shard_run_p2p(bsx_shard *s, const uint64_t *input, int32_t mem, int32_t input_is_coeffs, bsx_tree **tree_out,
                          uint64_t *rank_caps_host, bsx_timing *timing);
/* MerkleTree over this rank's slab (borrowed): 2^(cap_height - log2 G) local cap entries */
int32_t bsx_shard_commit(bsx_shard *s, bsx_tree **out);
/* this rank's polynomials in natural order, (c1 - c0) x 2^log_n */
int32_t bsx_shard_coeffs(bsx_shard *s, uint64_t *out, int32_t mem);

/* ---- the same, driven by ONE process over the GPUs of a box (SURVEY.md 8b bsx_comm_create) -------------------
 * A communicator owns one bsx_ctx per rank (devices[r]; n_ranks a power of two <= BSX_MAX_RANKS; a device may appear
 * more than once, the ranks then share it).  bsx_sharded_from_values runs the call order above with one host thread
 * per rank and peer pointers between the devices (NVLink peer access); no collective library is involved.  The result
 * is word for word the single-GPU PolynomialBatch: same cap, same rows, same Merkle paths, same FRI proof. */
typedef struct bsx_comm bsx_comm;
typedef struct bsx_sharded_batch bsx_sharded_batch;
int32_t bsx_comm_create(int32_t n_ranks, const int32_t *devices, bsx_comm **out);
void bsx_comm_destroy(bsx_comm *comm);
int32_t bsx_comm_size(const bsx_comm *comm);
bsx_ctx *bsx_comm_ctx(bsx_comm *comm, int32_t rank);
const char *bsx_comm_last_error(const bsx_comm *comm);
/* PolynomialBatch::from_values (or from_coeffs) over the communicator.  values_host: n_cols x 2^log_n, column-major,
 * host memory (pinned or pageable).  timing: the slowest rank of every phase. */
int32_t bsx_sharded_from_values(bsx_comm *comm, const uint64_t *values_host, int32_t log_n, int32_t n_cols,
                                int32_t rate_bits, int32_t cap_height, int32_t input_is_coeffs, bsx_sharded_batch **out,
                                bsx_timing *timing);
/* the same with one pointer per rank to that rank's columns (bsx_shard_column_range), host memory or memory of that
 * rank's device (`mem`) */
int32_t bsx_sharded_from_values_parts(bsx_comm *comm, const uint64_t *const *parts, int32_t mem, int32_t log_n,
                                      int32_t n_cols, int32_t rate_bits, int32_t cap_height, int32_t input_is_coeffs,
                                      bsx_sharded_batch **out, bsx_timing *timing);
void bsx_sharded_free(bsx_sharded_batch *b);
int32_t bsx_sharded_cap(const bsx_sharded_batch *b, uint64_t *out_host /* 2^cap_height x 4 */);
/* MerkleTree::get / prove with GLOBAL leaf indices, answered by the owning rank; the proof verifies against the cap of
 * the whole batch and has bsx_sharded_proof_len() = log2(leaves) - cap_height siblings */
int32_t bsx_sharded_get(bsx_sharded_batch *b, const uint64_t *leaf_idx_host, int32_t k, uint64_t *out_host);
int32_t bsx_sharded_prove(bsx_sharded_batch *b, uint64_t leaf_index, uint64_t *siblings_host);
int32_t bsx_sharded_proof_len(const bsx_sharded_batch *b);
/* this rank's part, for the lower-level calls (bsx_shard_coeffs, bsx_merkle_*) */
bsx_shard *bsx_sharded_shard(bsx_sharded_batch *b, int32_t rank);
bsx_tree *bsx_sharded_tree(bsx_sharded_batch *b, int32_t rank);

/* ---- Challenger (iop/challenger.rs) and FRI (fri/oracle.rs prove_openings, fri/prover.rs) -------------
 * bsx_challenger mirrors plonky2's `Challenger<F, PoseidonHash>`: sponge state, input buffer, output
 * buffer (outputs are popped from the BACK, as Vec::pop does).  The struct lives on the host between calls;
 * the duplexing permutations run on the device. */
typedef struct bsx_challenger {
    uint64_t sponge_state[12];
    uint64_t input_buffer[8];
    uint64_t output_buffer[8];
    int32_t n_input, n_output;
} bsx_challenger;
/* Challenger::new() */
void bsx_challenger_init(bsx_challenger *ch);
/* Challenger::observe_elements / observe_cap / observe_hash (n words) */
int32_t bsx_challenger_observe(bsx_ctx *ctx, bsx_challenger *ch, const uint64_t *elems_host, size_t n);
/* Challenger::get_n_challenges (get_extension_challenge = 2 challenges as [c0, c1]) */
int32_t bsx_challenger_get(bsx_ctx *ctx, bsx_challenger *ch, uint64_t *out_host, int32_t n);

/* FriInstanceInfo.batches: per opening batch a point (extension) and its (oracle index, polynomial index) list */
typedef struct bsx_fri_instance {
    int32_t n_batches;
    const uint64_t *points;     /* n_batches x 2 */
    const int32_t *n_polys;     /* per batch */
    const int32_t *poly_oracle; /* concatenated over batches */
    const int32_t *poly_index;
} bsx_fri_instance;
/* FriParams + FriConfig (fri/mod.rs); reduction_arity_bits as produced by FriReductionStrategy */
typedef struct bsx_fri_params {
    int32_t degree_bits, rate_bits, cap_height, proof_of_work_bits, num_query_rounds;
    int32_t n_arity;
    int32_t arity_bits[16];
} bsx_fri_params;
/* Words of the flat FriProof stream for these oracles/params.  Layout, in plonky2 `FriProof` field order:
 *   commit_phase_merkle_caps[n_arity][2^cap_height][4]
 *   query_round_proofs[num_query_rounds]: initial_trees_proof per oracle { leaf[width], siblings[..][4] },
 *                                         steps per round { evals[2^arity][2], siblings[..][4] }
 *   final_poly[2^(degree_bits - sum arity)][2]
 *   pow_witness */
int32_t bsx_fri_proof_words(bsx_batch *const *oracles, int32_t n_oracles, const bsx_fri_params *params, size_t *out);
/* Replaces PolynomialBatch::prove_openings(instance, oracles, challenger, fri_params): draws alpha, builds
 * the final polynomial (alpha-combination, division by (X - z), times X), LDE, then fri_proof.
 * The proof-of-work witness is the SMALLEST valid one (upstream takes any).
 * challenger is advanced exactly as upstream's is.  proof_host receives bsx_fri_proof_words() words. */
int32_t bsx_fri_prove_openings(bsx_ctx *ctx, bsx_batch *const *oracles, int32_t n_oracles,
                               const bsx_fri_instance *instance, bsx_challenger *challenger,
                               const bsx_fri_params *params, uint64_t *proof_host);
/* Replaces fri::prover::fri_proof(initial_merkle_trees, lde_polynomial_coeffs, lde_polynomial_values, ..):
 * lde_coeffs_host = 2^(degree_bits + rate_bits) extension coefficients [c0, c1] (only the first 1/rate are
 * non-zero); the values are recomputed on the device. */
int32_t bsx_fri_proof(bsx_ctx *ctx, bsx_batch *const *oracles, int32_t n_oracles, const uint64_t *lde_coeffs_host,
                      bsx_challenger *challenger, const bsx_fri_params *params, uint64_t *proof_host);
/* ---- prove_openings over row-sharded oracles (FRI after gathering the small final-poly inputs) ----
 * 1. every rank: alpha = challenger.get_extension_challenge() (identical transcripts), then its part of the final
 *    polynomial from the coefficient columns it holds: bsx_shard_final_poly_part (or bsx_fri_final_poly_part on
 *    whole batches, with poly_oracle = -1 for polynomials held elsewhere -- they keep their alpha power);
 * 2. the parts are gathered (16 bytes x 2^degree_bits per rank) and added: bsx_field_sum;
 * 3. bsx_fri_proof_from_final_poly runs the commit phase, proof of work and the FRI-tree query openings and returns
 *    the query indices; with n_oracles = 0 the proof stream has no initial-tree openings;
 * 4. the owning ranks serve the initial-tree openings of those indices (bsx_merkle_get / bsx_merkle_prove on their
 *    row slabs) and the host splices them in (0-kno-blobstreamx_b200/serialization.py splice_initial_openings). */
int32_t bsx_fri_final_poly_part(bsx_ctx *ctx, bsx_batch *const *oracles, int32_t n_oracles,
                                const bsx_fri_instance *instance, const uint64_t alpha[2], int32_t degree_bits,
                                uint64_t *out, int32_t mem_out);
int32_t bsx_shard_final_poly_part(bsx_shard *const *shards, int32_t n_shards, const bsx_fri_instance *instance,
                                  const uint64_t alpha[2], uint64_t *out, int32_t mem_out);
/* out[i] = sum over k vectors of `words` words (mod p) */
int32_t bsx_field_sum(bsx_ctx *ctx, const uint64_t *parts, int32_t mem, int32_t k, size_t words, uint64_t *out,
                      int32_t mem_out);
/* final_poly: 2^degree_bits x [c0, c1] on the `mem` side; query_indices_host: num_query_rounds words or NULL */
int32_t bsx_fri_proof_from_final_poly(bsx_ctx *ctx, bsx_batch *const *oracles, int32_t n_oracles,
                                      const uint64_t *final_poly, int32_t mem, bsx_challenger *challenger,
                                      const bsx_fri_params *params, uint64_t *proof_host, uint64_t *query_indices_host);
/* test hook: the final polynomial of prove_openings for a given alpha -> out_host[2^degree_bits][2] */
int32_t bsx_fri_final_poly(bsx_ctx *ctx, bsx_batch *const *oracles, int32_t n_oracles,
                           const bsx_fri_instance *instance, const uint64_t alpha[2], uint64_t *out_host);

/* PolynomialBatch::prove_openings over sharded oracles in one call (steps 1-4 above inside one process): the proof
 * equals bsx_fri_prove_openings on the unsharded batches word for word.  `instance` uses GLOBAL polynomial indices. */
int32_t bsx_sharded_prove_openings(bsx_comm *comm, bsx_sharded_batch *const *oracles, int32_t n_oracles,
                                   const bsx_fri_instance *instance, bsx_challenger *challenger,
                                   const bsx_fri_params *params, uint64_t *proof_host);

/* ---- typed view of the flat FriProof stream, and proof bytes (plonk/proof.rs, util/serialization/mod.rs) -------
 * bsx_fri_proof_layout: word offsets of every plonky2 `FriProof` field inside the stream of bsx_fri_proof_words, so a
 * caller (the Rust shim building `FriProof { commit_phase_merkle_caps, query_round_proofs, final_poly, pow_witness }`)
 * can slice it without knowing the packing.  Offsets inside a query round are relative to
 * query_offset + q * query_stride.  leaf_widths[o] = polynomials + salt columns of oracle o. */
typedef struct bsx_fri_layout {
    uint64_t total_words;
    uint64_t caps_offset, cap_words;          /* commit_phase_merkle_caps[r] at caps_offset + r * cap_words */
    uint64_t query_offset, query_stride;      /* query_round_proofs[q] */
    uint64_t initial_leaf_offset[BSX_MAX_ORACLES], initial_leaf_words[BSX_MAX_ORACLES]; /* initial_trees_proof.evals_proofs[o].0 */
    uint64_t initial_proof_offset[BSX_MAX_ORACLES];                                      /* .1.siblings, initial_siblings x 4 */
    uint64_t step_evals_offset[16], step_evals_words[16];                               /* steps[r].evals, [c0, c1] pairs */
    uint64_t step_proof_offset[16];                                                      /* steps[r].merkle_proof.siblings */
    uint64_t final_poly_offset, final_poly_words; /* final_poly.coeffs, [c0, c1] pairs */
    uint64_t pow_witness_offset;
    int32_t n_oracles, n_rounds, n_query_rounds, initial_siblings;
    int32_t step_siblings[16];
} bsx_fri_layout;
int32_t bsx_fri_proof_layout(const bsx_fri_params *params, const int32_t *leaf_widths, int32_t n_oracles,
                             bsx_fri_layout *out);
/* `Write::write_fri_proof` of the stream.  out == NULL only counts (out_len); BSX_ERR_INVALID_ARG when out_cap is short. */
int32_t bsx_fri_proof_bytes(const uint64_t *proof_words, const bsx_fri_params *params, const int32_t *leaf_widths,
                            int32_t n_oracles, uint8_t *out, size_t out_cap, size_t *out_len);
/* `ProofWithPublicInputs::to_bytes()`: Proof { wires_cap, plonk_zs_partial_products_cap, quotient_polys_cap, openings,
 * opening_proof } then the public inputs.  openings[k]: n_openings[k] extension elements [c0, c1] in `OpeningSet` field
 * order (BSX_OPEN_*).  All pointers are host memory. */
#define BSX_OPENING_SET_FIELDS 7
#define BSX_OPEN_CONSTANTS 0
#define BSX_OPEN_PLONK_SIGMAS 1
#define BSX_OPEN_WIRES 2
#define BSX_OPEN_PLONK_ZS 3
#define BSX_OPEN_PLONK_ZS_NEXT 4
#define BSX_OPEN_PARTIAL_PRODUCTS 5
#define BSX_OPEN_QUOTIENT_POLYS 6
typedef struct bsx_proof_parts {
    const uint64_t *wires_cap, *plonk_zs_partial_products_cap, *quotient_polys_cap; /* 2^cap_height x 4 each */
    const uint64_t *openings[BSX_OPENING_SET_FIELDS];
    int32_t n_openings[BSX_OPENING_SET_FIELDS];
    const uint64_t *opening_proof_words; /* stream of bsx_fri_prove_openings */
    const uint64_t *public_inputs;
    int32_t n_public_inputs;
} bsx_proof_parts;
int32_t bsx_proof_with_public_inputs_bytes(const bsx_proof_parts *parts, const bsx_fri_params *params,
                                           const int32_t *leaf_widths, int32_t n_oracles, uint8_t *out, size_t out_cap,
                                           size_t *out_len);

/* Replaces the `eval_commitment` closure of plonk/proof.rs OpeningSet::new: every polynomial
 * col_begin .. col_begin + n_cols - 1 of the batch (salt columns are not polynomials) evaluated at n_points
 * extension-field points, `PolynomialCoeffs::to_extension().eval(z)`.  points_host = n_points x [c0, c1];
 * out_host = [n_points][n_cols][2].  The coefficients are read from HBM once for all points. */
int32_t bsx_batch_eval_ext(bsx_batch *batch, int32_t col_begin, int32_t n_cols, const uint64_t *points_host,
                           int32_t n_points, uint64_t *out_host);

/* Replaces the hand-over from quotient evaluation to the quotient commitment in plonk/prover.rs: the tail of
 * compute_quotient_polys (`PolynomialValues::coset_ifft(F::coset_shift())` of every challenge's quotient values) and,
 * in prove(), `quotient_poly.trim_to_len(quotient_degree)`, `.chunks(degree)` and
 * `PolynomialBatch::from_coeffs(all_quotient_poly_chunks, rate_bits, blinding, cap_height, ..)`.
 * values = n_polys x 2^(degree_log + quotient_degree_bits) words, column-major, natural order: the quotient's values on
 * the coset 7 * <omega_{N 2^q}>.  quotient_degree_factor <= 2^quotient_degree_bits chunks of 2^degree_log coefficients are
 * kept per polynomial; non-zero coefficients beyond them fail with BSX_ERR_INVALID_ARG (upstream `trim_to_len` returns
 * Err and prove() panics "Quotient has failed ...").  The batch has n_polys * quotient_degree_factor polynomials.
 * `salt` (optional, with `blinding`) lives on the same side as `values` (`mem`), laid out as for bsx_batch_from_coeffs. */
int32_t bsx_batch_from_quotient_values(bsx_ctx *ctx, const uint64_t *values, int32_t mem, int32_t degree_log,
                                       int32_t quotient_degree_bits, int32_t quotient_degree_factor, int32_t n_polys,
                                       int32_t rate_bits, int32_t cap_height, int32_t blinding, const uint64_t *salt,
                                       bsx_batch **out, bsx_timing *timing);

/* Replaces plonk/prover.rs `all_wires_permutation_partial_products` (per challenge
 * `wires_permutation_partial_products_and_zs`, with `quotient_chunk_products` and `partial_products_and_z_gx` of
 * plonk/permutation_argument.rs / util/partial_products.rs): the prover step between the wires commitment and the
 * Zs commitment.  wires, sigmas = n_routed x 2^log_n words, column-major, natural order: the routed wire values of
 * the witness and the values of the sigma polynomials on the subgroup (`prover_data.sigmas`, transposed).
 * Row i (x = omega^i), challenge c, chunk m of max_degree (= quotient_degree_factor) wires:
 *   q_m = prod_j (w_j + beta_c k_is[j] x + gamma_c) / prod_j (w_j + beta_c sigma_j + gamma_c);
 *   Z_c(1) = 1, Z_c(g x) = Z_c(x) prod_m q_m, partial product m = Z_c(x) q_0 .. q_m  (m < num_partial_products).
 * out = (n_challenges * ceil(n_routed / max_degree)) x 2^log_n values in the column order of the commitment that
 * follows ("Z is expected at the front of our batch"): Z_0 .. Z_{n_challenges - 1}, then the num_partial_products
 * columns of challenge 0, of challenge 1, ...  -- ready for bsx_batch_from_values.  A zero denominator fails with
 * BSX_ERR_INVALID_ARG (upstream panics in batch_multiplicative_inverse). */
int32_t bsx_partial_products_and_zs(bsx_ctx *ctx, const uint64_t *wires, const uint64_t *sigmas, int32_t mem,
                                    int32_t log_n, int32_t n_routed, int32_t max_degree, const uint64_t *k_is_host,
                                    const uint64_t *betas_host, const uint64_t *gammas_host, int32_t n_challenges,
                                    uint64_t *out, int32_t out_mem);

/* ---- quotient evaluation on the device (plonk/prover.rs compute_quotient_polys) ------------------------------
 * The circuit as compute_quotient_polys and plonk/vanishing_poly.rs eval_vanishing_poly_base_batch read it from
 * CommonCircuitData / CircuitConfig.  Gates are data, not code: a table of (kind, parameter, selector info); the fourteen
 * kinds below are built in (the gates of plonky2's own recursion circuits; extension gates for D = 2); any other kind makes
 * the calls below fail with BSX_ERR_INVALID_ARG.  `row` is the gate's index
 * in common_data.gates (the value its selector polynomial takes on rows that use it), selector_index / group_* are
 * selectors_info.selector_indices[row] and selectors_info.groups[selector_index]. */
#define BSX_GATE_NOOP 0
#define BSX_GATE_CONSTANT 1     /* param = num_consts: constants[i] - wires[i] */
#define BSX_GATE_PUBLIC_INPUT 2 /* wires[i] - public_inputs_hash[i], i < 4 */
#define BSX_GATE_ARITHMETIC 3   /* param = num_ops: wires[4i+3] - (wires[4i] wires[4i+1] const_0 + wires[4i+2] const_1) */
#define BSX_GATE_POSEIDON 4     /* gates/poseidon.rs PoseidonGate: 135 wires, 123 constraints of degree 7 (swap binary, 4 deltas,
                                   S-box input wires of rounds 1..3 / the 22 partial rounds / the last 4 rounds, 12 outputs) */
#define BSX_GATE_ARITHMETIC_EXTENSION 5 /* gates/arithmetic_extension.rs, D = 2, param = num_ops: wires 8i .. 8i+7 =
                                   multiplicand_0, multiplicand_1, addend, output in F_p[X]/(X^2 - 7); 2 constraints each */
#define BSX_GATE_MUL_EXTENSION 6 /* gates/multiplication_extension.rs, D = 2, param = num_ops: wires 6i .. 6i+5 */
#define BSX_GATE_BASE_SUM 7     /* gates/base_sum.rs BaseSumGate<2>, param = num_limbs (<= 63): wire 0 = sum, wires 1.. = limbs */
#define BSX_GATE_EXPONENTIATION 8 /* gates/exponentiation.rs, param = num_power_bits: base 0, bits 1.., output 1 + n, intermediates 2 + n.. */
#define BSX_GATE_REDUCING 9     /* gates/reducing.rs ReducingGate<2>, param = num_coeffs: output, alpha, old_acc, coeffs, accumulators */
#define BSX_GATE_REDUCING_EXTENSION 10 /* gates/reducing_extension.rs ReducingExtensionGate<2>, param = num_coeffs */
#define BSX_GATE_RANDOM_ACCESS 11 /* gates/random_access.rs RandomAccessGate<2>: param = bits | num_copies << 8 |
                                   num_extra_constants << 16, bits <= 4 */
#define BSX_GATE_POSEIDON_MDS 12 /* gates/poseidon_mds.rs PoseidonMdsGate<2>: wires 24 + 2r + d = MDS row r of wires 2i + d */
#define BSX_GATE_COSET_INTERPOLATION 13 /* gates/coset_interpolation.rs CosetInterpolationGate<2>: param = subgroup_bits |
                                   degree << 8; at most one per circuit */
typedef struct bsx_gate {
    int32_t kind, param, row, selector_index, group_start, group_end;
} bsx_gate;
typedef struct bsx_circuit {
    int32_t degree_bits, rate_bits, quotient_degree_factor, num_challenges;
    int32_t num_wires, num_routed_wires;
    int32_t num_constants; /* common_data.num_constants: selector columns first, then the gates' constants */
    int32_t num_selectors;
    int32_t num_partial_products, num_gate_constraints, n_gates;
    const bsx_gate *gates;  /* host */
    const uint64_t *k_is;   /* host, num_routed_wires */
} bsx_circuit;
/* The quotient's values on the coset 7 <omega_{N 2^q}> (q = log2_ceil(quotient_degree_factor)) before coset_ifft:
 * out[a][i], a < num_challenges, natural order -- `quotient_values` of compute_quotient_polys, transposed.  The three
 * commitments stay in HBM: constants_sigmas (num_constants + num_routed_wires polynomials), wires (num_wires), Zs and
 * partial products (num_challenges * (1 + num_partial_products), Zs first), all of 2^degree_bits rows at rate_bits. */
int32_t bsx_quotient_values(bsx_ctx *ctx, const bsx_circuit *circuit, bsx_batch *constants_sigmas, bsx_batch *wires,
                            bsx_batch *zs_partial_products, const uint64_t *public_inputs_hash /* 4, host */,
                            const uint64_t *betas, const uint64_t *gammas, const uint64_t *alphas /* host */,
                            uint64_t *out, int32_t out_mem);
/* compute_quotient_polys and the quotient commitment of prove() in one call: the values above, coset_ifft,
 * trim_to_len(quotient_degree_factor * N) (fails where upstream panics "Quotient has failed"), chunks of N,
 * PolynomialBatch::from_coeffs.  eval_ms (optional): device milliseconds of the evaluation kernel alone. */
int32_t bsx_compute_quotient_polys(bsx_ctx *ctx, const bsx_circuit *circuit, bsx_batch *constants_sigmas, bsx_batch *wires,
                                   bsx_batch *zs_partial_products, const uint64_t *public_inputs_hash,
                                   const uint64_t *betas, const uint64_t *gammas, const uint64_t *alphas,
                                   int32_t cap_height, int32_t blinding, bsx_batch **out, bsx_timing *timing,
                                   float *eval_ms);

/* ---- test hooks -----------------------------------------------------------------------------
 * Poseidon::poseidon on n states (n x 12 words, in place). */
int32_t bsx_poseidon_permute(bsx_ctx *ctx, uint64_t *states, int32_t mem, size_t n);
/* gates/poseidon.rs PoseidonGenerator::run_once on n row-major rows of 135 wires (in place): the inputs (wires 0..11) and
 * the swap wire (24, 0 or 1) are read; the delta wires, every S-box input wire and the outputs are written.  Witness
 * generation itself is outside the path this library replaces; this fills the PoseidonGate rows of synthetic circuits
 * and checks bsx_quotient_values' constraints from the other side. */
int32_t bsx_poseidon_gate_witness(bsx_ctx *ctx, uint64_t *rows, int32_t mem, size_t n);
/* n words of the ChaCha20 (RFC 8439 block function, 64-bit block counter) key stream that the blinding salts are
 * drawn from under `key`; raw != 0 returns the stream itself (eight little-endian u64 per block, block t at words
 * 8t .. 8t+7), raw == 0 the salt words (candidates >= p rejected and redrawn).  In a commit the key is 256 bits from
 * getrandom(2) per batch and never leaves the call. */
int32_t bsx_debug_salt_stream(bsx_ctx *ctx, const uint32_t key[8], int32_t raw, uint64_t *out_host, size_t n);
/* Batched transforms on n_cols columns of 2^log_n words, column-major, in place, natural order in and
 * out.  kind: 0 = fft (PolynomialCoeffs::fft), 1 = ifft (PolynomialValues::ifft),
 * 2 = coset_fft(shift) (PolynomialCoeffs::coset_fft), 3 = coset_ifft(shift) (PolynomialValues::coset_ifft). */
int32_t bsx_ntt(bsx_ctx *ctx, uint64_t *data, int32_t mem, int32_t log_n, int32_t n_cols, int32_t kind,
                uint64_t shift);

#ifdef __cplusplus
}
#endif
#endif
