/* plonky2_oracle.h -- CPU restatement of the plonky2 commitment phase.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the shipped product path may include,
 * link or call this file: only tests/, __graft_entry__.smoke() and the
 * cpu_baseline / --impl reference legs of bench.py use it, and only as the checker
 * or as the timed CPU arm.
 *
 * PARITY STATUS: "parity unpinned" against the reference's own prover.  The
 * reference mount (/root/reference) holds no source and no
 * Rust toolchain exists here, so this file restates the PUBLIC plonky2 algorithm
 *  from the upstream files named beside each function.
 * What IS pinned: the Poseidon permutation against plonky2's published
 * known-answer vectors (hash/poseidon_goldilocks.rs `check_test_vectors`),
 * field constants, and NTT/Merkle/FRI self-consistency (naive evaluation,
 * proof verification, a FRI verifier restatement).
 *
 * All field elements are canonical little-endian u64 (< p = 2^64 - 2^32 + 1).
 * Extension elements F_p[x]/(x^2-7) are two consecutive u64 [c0, c1].
 */
#ifndef PLONKY2_ORACLE_H
#define PLONKY2_ORACLE_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define ORC_P 0xFFFFFFFF00000001ULL

/* ---- field (plonky2 field/src/goldilocks_field.rs, extension/quadratic.rs) ---- */
uint64_t orc_add(uint64_t a, uint64_t b);
uint64_t orc_sub(uint64_t a, uint64_t b);
uint64_t orc_mul(uint64_t a, uint64_t b);
uint64_t orc_pow(uint64_t a, uint64_t e);
uint64_t orc_inv(uint64_t a);
uint64_t orc_root_of_unity(int log_n);          /* primitive 2^log_n-th root */
void orc_ext_mul(const uint64_t a[2], const uint64_t b[2], uint64_t out[2]);
void orc_ext_inv(const uint64_t a[2], uint64_t out[2]);

/* ---- FFT family (field/src/fft.rs, field/src/polynomial/mod.rs) ---- */
void orc_fft(uint64_t *a, int log_n);                       /* natural in, natural out */
void orc_ifft(uint64_t *a, int log_n);
void orc_coset_fft(uint64_t *a, int log_n, uint64_t shift); /* a_j *= shift^j ; fft */
void orc_coset_ifft(uint64_t *a, int log_n, uint64_t shift); /* ifft ; a_j *= shift^-j */
/* PolynomialCoeffs::lde(rate_bits).coset_fft(shift): out has 2^(log_n+rate_bits) entries */
void orc_lde(const uint64_t *coeffs, int log_n, int rate_bits, uint64_t shift, uint64_t *out);
/* strided variants for extension polynomials stored interleaved [c0,c1,...] */
void orc_fft_strided(uint64_t *a, int log_n, size_t stride);
uint64_t orc_reverse_bits(uint64_t x, int bits);

/* ---- Poseidon (hash/poseidon.rs, poseidon_goldilocks.rs, hashing.rs) ---- */
void orc_poseidon(uint64_t state[12]);
void orc_hash_no_pad(const uint64_t *in, size_t len, uint64_t out[4]);
void orc_hash_or_noop(const uint64_t *in, size_t len, uint64_t out[4]);
void orc_two_to_one(const uint64_t l[4], const uint64_t r[4], uint64_t out[4]);
const uint64_t *orc_round_constants(void);
/* the two layers of one round, for the PoseidonGate constraints (canonical in, canonical out) */
uint64_t orc_poseidon_sbox(uint64_t x);
void orc_poseidon_mds(uint64_t state[12]);

/* ---- Merkle tree (hash/merkle_tree.rs, merkle_proofs.rs) ----
 * leaves: row-major n_leaves x leaf_len.  digests (level order, leaf level first):
 * level k has (n_leaves >> k) digests of 4 words, for k = 0 .. log_leaves-cap_height-1.
 * Total words = 4 * 2 * (n_leaves - 2^cap_height).  cap: 2^cap_height x 4 words.
 * Returns 0, or -1 on invalid arguments (upstream panics). */
int orc_merkle_build(const uint64_t *leaves, int log_leaves, size_t leaf_len, int cap_height,
                     uint64_t *digests, uint64_t *cap);
/* siblings bottom-up, (log_leaves - cap_height) x 4 words */
int orc_merkle_prove(const uint64_t *digests, int log_leaves, int cap_height, size_t leaf_index,
                     uint64_t *siblings);
/* 1 if the proof verifies against cap, else 0 */
int orc_merkle_verify(const uint64_t *leaf, size_t leaf_len, size_t leaf_index,
                      const uint64_t *siblings, int n_siblings, const uint64_t *cap, int cap_height);
/* level-order digests -> plonky2's recursive `digests` Vec layout (merkle_tree.rs fill_subtree) */
void orc_merkle_digests_to_plonky2(const uint64_t *level_digests, int log_leaves, int cap_height,
                                   uint64_t *out);

/* ---- PolynomialBatch::from_values / from_coeffs (fri/oracle.rs) ----
 * values/coeffs: column-major n_cols x 2^log_n.  salt (optional, blinding): 4 x 2^(log_n+rate_bits)
 * column-major LDE-domain salt columns appended after the real columns (upstream draws them at random).
 * leaves: row-major 2^(log_n+rate_bits) x (n_cols + n_salt), rows in bit-reversed order.
 * Any output pointer except cap may be NULL.  */
int orc_batch_from_values(const uint64_t *values, int log_n, int n_cols, int rate_bits,
                          int cap_height, const uint64_t *salt, int n_salt, uint64_t *coeffs,
                          uint64_t *leaves, uint64_t *digests, uint64_t *cap);
int orc_batch_from_coeffs(const uint64_t *coeffs, int log_n, int n_cols, int rate_bits,
                          int cap_height, const uint64_t *salt, int n_salt, uint64_t *leaves,
                          uint64_t *digests, uint64_t *cap);


/* ---- Challenger (iop/challenger.rs), Poseidon duplex sponge in overwrite mode ---- */
typedef struct orc_challenger {
    uint64_t state[12];
    uint64_t in[8];
    uint64_t out[8];
    int32_t n_in, n_out;
} orc_challenger;
void orc_challenger_init(orc_challenger *c);
void orc_challenger_observe(orc_challenger *c, uint64_t x);
void orc_challenger_observe_many(orc_challenger *c, const uint64_t *x, size_t n);
uint64_t orc_challenger_get(orc_challenger *c);
void orc_challenger_get_ext(orc_challenger *c, uint64_t out[2]);

/* ---- FRI (fri/oracle.rs prove_openings, fri/prover.rs, fri/verifier.rs) ---- */
typedef struct orc_batch_view { /* one committed PolynomialBatch as produced by orc_batch_from_* */
    int32_t log_n, n_cols, n_salt, rate_bits, cap_height;
    const uint64_t *coeffs;  /* n_cols x 2^log_n, natural order */
    const uint64_t *leaves;  /* 2^(log_n+rate_bits) x (n_cols+n_salt) row-major */
    const uint64_t *digests; /* level order */
    const uint64_t *cap;
} orc_batch_view;
typedef struct orc_fri_instance { /* FriInstanceInfo.batches: opening point + (oracle, polynomial) list each */
    int32_t n_batches;
    const uint64_t *points;     /* n_batches x 2 (extension) */
    const int32_t *n_polys;     /* per batch */
    const int32_t *poly_oracle; /* concatenated over batches */
    const int32_t *poly_index;
} orc_fri_instance;
typedef struct orc_fri_params { /* FriParams + FriConfig */
    int32_t degree_bits, rate_bits, cap_height, proof_of_work_bits, num_query_rounds;
    int32_t n_arity;
    int32_t arity_bits[16];
} orc_fri_params;
int orc_fri_final_poly(const orc_batch_view *oracles, int n_oracles, const orc_fri_instance *inst,
                       const uint64_t alpha[2], uint64_t *out /* 2^log_n x 2 */);
size_t orc_fri_proof_words(const orc_batch_view *oracles, int n_oracles, const orc_fri_params *fp);
/* PolynomialBatch::prove_openings: draws alpha from the challenger, builds the final polynomial,
 * LDE, then fri_proof (commit phase, PoW with the SMALLEST valid witness, query rounds).
 * proof: flat u64 stream, layout in plonky2_fri_oracle.c.  The challenger is advanced exactly as upstream. */
int orc_fri_prove_openings(const orc_batch_view *oracles, int n_oracles, const orc_fri_instance *inst,
                           orc_challenger *ch, const orc_fri_params *fp, uint64_t *proof);
/* fri/verifier.rs verify_fri_proof restated.  openings: per opening batch, the claimed evaluations of
 * its polynomials (extension, concatenated in instance order).  `ch` must be in the state the prover's
 * challenger had when prove_openings was called.  Returns 1 when the proof verifies, 0 otherwise. */
int orc_fri_verify(const orc_batch_view *oracles /* only shapes + cap used */, int n_oracles,
                   const orc_fri_instance *inst, const uint64_t *openings, orc_challenger *ch,
                   const orc_fri_params *fp, const uint64_t *proof);
/* evaluate polynomial `poly_index` of an oracle (coefficient form) at an extension point */
void orc_eval_poly_ext(const orc_batch_view *o, int poly_index, const uint64_t z[2], uint64_t out[2]);

int orc_num_threads(void);
void orc_set_num_threads(int n);

/* plonk/prover.rs wires_permutation_partial_products_and_zs for one challenge; -1 on a zero denominator */
int orc_partial_products_and_z(const uint64_t *wires, const uint64_t *sigmas, int log_n, int n_routed, int max_degree,
                               const uint64_t *k_is, uint64_t beta, uint64_t gamma, uint64_t *out_z, uint64_t *out_pp);

/* ---- quotient evaluation (plonky2_plonk_oracle.c): plonk/prover.rs compute_quotient_polys,
 * plonk/vanishing_poly.rs eval_vanishing_poly_base_batch, the gates of gates/ ---- */
#define ORC_GATE_NOOP 0         /* gates/noop.rs NoopGate */
#define ORC_GATE_CONSTANT 1     /* gates/constant.rs ConstantGate { num_consts = param } */
#define ORC_GATE_PUBLIC_INPUT 2 /* gates/public_input.rs PublicInputGate */
#define ORC_GATE_ARITHMETIC 3   /* gates/arithmetic_base.rs ArithmeticGate { num_ops = param } */
#define ORC_GATE_POSEIDON 4     /* gates/poseidon.rs PoseidonGate: 135 wires, 123 constraints of degree 7 */
#define ORC_GATE_ARITHMETIC_EXTENSION 5 /* gates/arithmetic_extension.rs ArithmeticExtensionGate<2> { num_ops = param } */
#define ORC_GATE_MUL_EXTENSION 6        /* gates/multiplication_extension.rs MulExtensionGate<2> { num_ops = param } */
#define ORC_GATE_BASE_SUM 7             /* gates/base_sum.rs BaseSumGate<2> { num_limbs = param } */
#define ORC_GATE_EXPONENTIATION 8       /* gates/exponentiation.rs ExponentiationGate { num_power_bits = param } */
#define ORC_GATE_REDUCING 9             /* gates/reducing.rs ReducingGate<2> { num_coeffs = param } */
#define ORC_GATE_REDUCING_EXTENSION 10  /* gates/reducing_extension.rs ReducingExtensionGate<2> { num_coeffs = param } */
#define ORC_GATE_RANDOM_ACCESS 11       /* gates/random_access.rs RandomAccessGate<2>: param = bits | num_copies << 8 |
                                           num_extra_constants << 16 */
#define ORC_GATE_POSEIDON_MDS 12        /* gates/poseidon_mds.rs PoseidonMdsGate<2>: 24 inputs -> 24 outputs */
#define ORC_GATE_COSET_INTERPOLATION 13 /* gates/coset_interpolation.rs CosetInterpolationGate<2>: param = subgroup_bits |
                                           degree << 8 */
typedef struct orc_gate { /* one entry of common_data.gates with its share of common_data.selectors_info */
    int32_t kind, param;
    int32_t row;                    /* index in common_data.gates: the selector value that switches it on */
    int32_t selector_index;         /* selectors_info.selector_indices[row] */
    int32_t group_start, group_end; /* selectors_info.groups[selector_index] */
} orc_gate;
typedef struct orc_circuit { /* the fields of CommonCircuitData / CircuitConfig compute_quotient_polys reads */
    int32_t degree_bits, rate_bits, quotient_degree_factor, num_challenges;
    int32_t num_wires, num_routed_wires;
    int32_t num_constants; /* common_data.num_constants: selectors first, then the gates' constants */
    int32_t num_selectors; /* selectors_info.num_selectors() */
    int32_t num_partial_products, num_gate_constraints, n_gates;
    const orc_gate *gates;
    const uint64_t *k_is;
} orc_circuit;
int orc_gate_num_constraints(const orc_gate *g);
/* gates/poseidon.rs PoseidonGenerator: fills deltas, S-box input wires and outputs of n row-major rows of 135 wires */
void orc_poseidon_gate_fill(uint64_t *rows, size_t n);
/* the quotient values on the coset before `coset_ifft`: out[a][i], a < num_challenges, i < 2^(degree_bits + q) */
int orc_quotient_values(const orc_circuit *c, const orc_batch_view *constants_sigmas, const orc_batch_view *wires_b,
                        const orc_batch_view *zs_b, const uint64_t public_inputs_hash[4], const uint64_t *betas,
                        const uint64_t *gammas, const uint64_t *alphas, uint64_t *out);
/* the vanishing polynomial at a base-field point from the polynomials' values there (verifier side: V(x) must equal
 * Z_H(x) * sum_i x^(n i) quotient_chunk_i(x)); res[num_challenges] */
int orc_vanishing_at(const orc_circuit *c, uint64_t x, const uint64_t *constants_sigmas_row, const uint64_t *wires_row,
                     const uint64_t *zs_pp_row, const uint64_t *next_zs_row, const uint64_t pi_hash[4],
                     const uint64_t *betas, const uint64_t *gammas, const uint64_t *alphas, uint64_t *res);
/* verifier side (plonk/vanishing_poly.rs eval_vanishing_poly): the reduced vanishing polynomial at an EXTENSION point from
 * the opening set; openings are [n][2] extension elements, res[num_challenges][2] */
int orc_vanishing_at_ext(const orc_circuit *c, const uint64_t zeta[2], const uint64_t *constants, const uint64_t *sigmas,
                         const uint64_t *wires, const uint64_t *zs, const uint64_t *zs_next, const uint64_t *partial_products,
                         const uint64_t pi_hash[4], const uint64_t *betas, const uint64_t *gammas, const uint64_t *alphas,
                         uint64_t *res);
/* util/serialization write_proof_with_public_inputs over the flat FriProof stream; returns the byte count */
size_t orc_write_proof_with_public_inputs(const uint64_t *wires_cap, const uint64_t *zs_cap, const uint64_t *quotient_cap,
                                          const uint64_t *const openings[7], const int n_openings[7],
                                          const uint64_t *fri_words, const orc_fri_params *fp, const int *widths,
                                          int n_oracles, const uint64_t *public_inputs, int n_public_inputs, uint8_t *out);

#ifdef __cplusplus
}
#endif
#endif
