//! Replaces, in plonky2/src/plonk/prover.rs: `all_wires_permutation_partial_products` (the step between the wires
//! and the Zs commitments), `compute_quotient_polys` with plonk/vanishing_poly.rs `eval_vanishing_poly_base_batch`,
//! and the quotient commitment of `prove()` (`trim_to_len`, `.chunks(degree)`, `PolynomialBatch::from_coeffs`).
//!
//! With these on the device the LDE of the wires (9 GB at 2^20 x 135) never crosses PCIe: upstream's CPU evaluator
//! reads every LDE row of the constants+sigmas, wires and Zs commitments.
//!
//! Gates are data for the library (a table of kind + parameter + selector info); the kinds it knows are NoopGate,
//! ConstantGate, PublicInputGate and ArithmeticGate.  A circuit with any other gate is refused here (no fallback in
//! the library; the caller keeps upstream's CPU evaluator for such circuits and pays the LDE download).
use std::ptr::NonNull;

use bsx_b200_sys as sys;
use plonky2::hash::hash_types::{HashOut, RichField};
use plonky2::plonk::circuit_data::CommonCircuitData;
use plonky2::plonk::config::{GenericConfig, Hasher};
use plonky2_field::extension::Extendable;
use plonky2_field::polynomial::PolynomialValues;
use plonky2_field::types::PrimeField64;
use plonky2_util::log2_strict;

use crate::oracle::PolynomialBatch;
use crate::{as_words, with_ctx, words_to_field};

/// `common_data.gates` + `common_data.selectors_info` as the table the library reads; None when a gate is not one of
/// the kinds the library evaluates (Noop, Constant, PublicInput, Arithmetic, Poseidon, PoseidonMds, ArithmeticExtension,
/// MulExtension, BaseSum<2>, Exponentiation, Reducing, ReducingExtension, RandomAccess up to 4 bits, CosetInterpolation).
pub fn gate_table<F: RichField + Extendable<D>, const D: usize>(common_data: &CommonCircuitData<F, D>) -> Option<Vec<sys::bsx_gate>> {
    let info = &common_data.selectors_info;
    common_data
        .gates
        .iter()
        .enumerate()
        .map(|(row, g)| {
            let id = g.0.id();
            let (kind, param) = if id.starts_with("NoopGate") {
                (sys::BSX_GATE_NOOP, 0)
            } else if id.starts_with("ConstantGate") {
                (sys::BSX_GATE_CONSTANT, g.0.num_constants() as i32)
            } else if id.starts_with("PublicInputGate") {
                (sys::BSX_GATE_PUBLIC_INPUT, 0)
            } else if id.starts_with("ArithmeticGate") {
                (sys::BSX_GATE_ARITHMETIC, g.0.num_constraints() as i32) // one constraint per operation
            } else if id.starts_with("PoseidonGate") && D == 2 {
                (sys::BSX_GATE_POSEIDON, 0)
            } else if id.starts_with("ArithmeticExtensionGate") && D == 2 {
                (sys::BSX_GATE_ARITHMETIC_EXTENSION, (g.0.num_constraints() / D) as i32) // D constraints per operation
            } else if id.starts_with("MulExtensionGate") && D == 2 {
                (sys::BSX_GATE_MUL_EXTENSION, (g.0.num_constraints() / D) as i32)
            } else if id.starts_with("BaseSumGate") && id.contains("Base: 2") {
                (sys::BSX_GATE_BASE_SUM, (g.0.num_constraints() - 1) as i32) // 1 + num_limbs constraints
            } else if id.starts_with("ExponentiationGate") {
                (sys::BSX_GATE_EXPONENTIATION, (g.0.num_constraints() - 1) as i32) // num_power_bits + 1 constraints
            } else if id.starts_with("ReducingExtensionGate") && D == 2 {
                (sys::BSX_GATE_REDUCING_EXTENSION, (g.0.num_constraints() / D) as i32)
            } else if id.starts_with("ReducingGate") && D == 2 {
                (sys::BSX_GATE_REDUCING, (g.0.num_constraints() / D) as i32)
            } else if id.starts_with("RandomAccessGate") && D == 2 {
                // "RandomAccessGate { bits: b, num_copies: c, num_extra_constants: e, .. }"
                let field = |name: &str| -> Option<i32> {
                    let at = id.find(name)? + name.len();
                    id[at..].trim_start_matches(|ch: char| ch == ':' || ch == ' ').split(|ch: char| !ch.is_ascii_digit()).next()?.parse().ok()
                };
                let (bits, copies, extra) = (field("bits")?, field("num_copies")?, field("num_extra_constants")?);
                if bits > 4 {
                    return None;
                }
                (sys::BSX_GATE_RANDOM_ACCESS, bits | copies << 8 | extra << 16)
            } else if id.starts_with("PoseidonMdsGate") && D == 2 {
                (sys::BSX_GATE_POSEIDON_MDS, 0)
            } else if id.starts_with("CosetInterpolationGate") && D == 2 {
                // "CosetInterpolationGate { subgroup_bits: b, degree: d, barycentric_weights: [..] }"; Gate::degree() is d
                let at = id.find("subgroup_bits")? + "subgroup_bits".len();
                let bits: i32 = id[at..].trim_start_matches(|ch: char| ch == ':' || ch == ' ').split(|ch: char| !ch.is_ascii_digit()).next()?.parse().ok()?;
                (sys::BSX_GATE_COSET_INTERPOLATION, bits | (g.0.degree() as i32) << 8)
            } else {
                return None;
            };
            let sel = info.selector_indices[row];
            let group = &info.groups[sel];
            Some(sys::bsx_gate { kind, param, row: row as i32, selector_index: sel as i32, group_start: group.start as i32, group_end: group.end as i32 })
        })
        .collect()
}

fn circuit<'a, F: RichField + Extendable<D>, const D: usize>(
    common_data: &CommonCircuitData<F, D>,
    gates: &'a [sys::bsx_gate],
    k_is: &'a [u64],
) -> sys::bsx_circuit {
    let cfg = &common_data.config;
    sys::bsx_circuit {
        degree_bits: common_data.degree_bits() as i32,
        rate_bits: cfg.fri_config.rate_bits as i32,
        quotient_degree_factor: common_data.quotient_degree_factor as i32,
        num_challenges: cfg.num_challenges as i32,
        num_wires: cfg.num_wires as i32,
        num_routed_wires: cfg.num_routed_wires as i32,
        num_constants: common_data.num_constants as i32,
        num_selectors: common_data.selectors_info.num_selectors() as i32,
        num_partial_products: common_data.num_partial_products as i32,
        num_gate_constraints: common_data.num_gate_constraints as i32,
        n_gates: gates.len() as i32,
        gates: gates.as_ptr(),
        k_is: k_is.as_ptr(),
    }
}

/// `all_wires_permutation_partial_products(witness, betas, gammas, prover_data, common_data)` followed by prove()'s
/// reordering: the value vectors of the Zs commitment, Zs first.  wires / sigmas: the routed columns' values on H.
pub fn partial_products_and_zs<F: RichField + Extendable<D>, const D: usize>(
    routed_wires: &[Vec<F>],
    sigmas: &[Vec<F>],
    betas: &[F],
    gammas: &[F],
    common_data: &CommonCircuitData<F, D>,
) -> Vec<PolynomialValues<F>> {
    let n = routed_wires[0].len();
    let log_n = log2_strict(n);
    let n_routed = routed_wires.len();
    let max_degree = common_data.quotient_degree_factor;
    let n_chunks = (n_routed + max_degree - 1) / max_degree;
    let flat = |cols: &[Vec<F>]| -> Vec<F> { cols.iter().flat_map(|c| c.iter().copied()).collect() };
    let (w, s) = (flat(routed_wires), flat(sigmas));
    let words = |v: &[F]| -> Vec<u64> { v.iter().map(|x| x.to_canonical_u64()).collect() };
    let (k_is, b, g) = (words(&common_data.k_is), words(betas), words(gammas));
    let mut out = vec![0u64; betas.len() * n_chunks * n];
    with_ctx(|c| {
        // a zero denominator fails like upstream's batch_multiplicative_inverse ("Tried to invert zero")
        c.check(unsafe {
            sys::bsx_partial_products_and_zs(c.raw(), as_words(&w), as_words(&s), sys::BSX_HOST, log_n as i32, n_routed as i32,
                                             max_degree as i32, k_is.as_ptr(), b.as_ptr(), g.as_ptr(), betas.len() as i32,
                                             out.as_mut_ptr(), sys::BSX_HOST)
        })
    });
    let all: Vec<F> = words_to_field(out);
    all.chunks(n).map(|c| PolynomialValues::new(c.to_vec())).collect()
}

/// `compute_quotient_polys(..)` + the quotient commitment of prove(): returns the committed quotient batch.  Panics
/// with upstream's message when the vanishing polynomial is not divisible by Z_H ("Quotient has failed").
#[allow(clippy::too_many_arguments)]
pub fn quotient_commitment<F: RichField + Extendable<D>, C: GenericConfig<D, F = F> + 'static, const D: usize>(
    common_data: &CommonCircuitData<F, D>,
    constants_sigmas_commitment: &PolynomialBatch<F, C, D>,
    wires_commitment: &PolynomialBatch<F, C, D>,
    zs_partial_products_commitment: &PolynomialBatch<F, C, D>,
    public_inputs_hash: &HashOut<F>,
    betas: &[F],
    gammas: &[F],
    alphas: &[F],
) -> PolynomialBatch<F, C, D>
where
    <C::Hasher as Hasher<F>>::Hash: From<[F; 4]>,
    C::Hasher: 'static,
{
    let gates = gate_table(common_data).expect("circuit uses a gate the device evaluator does not implement");
    let words = |v: &[F]| -> Vec<u64> { v.iter().map(|x| x.to_canonical_u64()).collect() };
    let k_is = words(&common_data.k_is);
    let circ = circuit(common_data, &gates, &k_is);
    let (pih, b, g, a) = (words(&public_inputs_hash.elements), words(betas), words(gammas), words(alphas));
    let cfg = &common_data.config;
    let mut raw = std::ptr::null_mut();
    with_ctx(|c| {
        c.check(unsafe {
            sys::bsx_compute_quotient_polys(c.raw(), &circ, constants_sigmas_commitment.raw.as_ptr(), wires_commitment.raw.as_ptr(),
                                            zs_partial_products_commitment.raw.as_ptr(), pih.as_ptr(), b.as_ptr(), g.as_ptr(),
                                            a.as_ptr(), cfg.fri_config.cap_height as i32, cfg.zero_knowledge as i32, &mut raw,
                                            std::ptr::null_mut(), std::ptr::null_mut())
        })
    });
    PolynomialBatch::from_raw(NonNull::new(raw).unwrap(), cfg.zero_knowledge)
}
