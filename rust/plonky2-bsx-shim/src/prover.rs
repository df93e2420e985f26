//! Replaces plonky2/src/fri/prover.rs `fri_proof` and fri/oracle.rs `PolynomialBatch::prove_openings`.
//!
//! The device returns the proof as one flat word stream in `FriProof` field order; `bsx_fri_proof_layout` gives the
//! offset of every field, and `fri_proof_from_words` rebuilds plonky2's struct from it.  The proof-of-work witness is
//! the SMALLEST valid one (upstream's rayon `find_any` takes any: its output is not reproducible either).
use bsx_b200_sys as sys;
use plonky2::fri::proof::{FriInitialTreeProof, FriProof, FriQueryRound, FriQueryStep};
use plonky2::fri::structure::FriInstanceInfo;
use plonky2::fri::FriParams;
use plonky2::hash::hash_types::RichField;
use plonky2::hash::merkle_proofs::MerkleProof;
use plonky2::hash::merkle_tree::MerkleCap;
use plonky2::iop::challenger::Challenger;
use plonky2::plonk::config::{GenericConfig, Hasher};
use plonky2::util::timing::TimingTree;
use plonky2_field::extension::{Extendable, FieldExtension};
use plonky2_field::polynomial::PolynomialCoeffs;
use plonky2_field::types::{Field, PrimeField64};

use crate::oracle::PolynomialBatch;
use crate::{hash_from_words, with_ctx};

fn fri_params(p: &FriParams) -> sys::bsx_fri_params {
    assert!(p.reduction_arity_bits.len() <= 16);
    let mut arity_bits = [0i32; 16];
    for (a, &b) in arity_bits.iter_mut().zip(&p.reduction_arity_bits) {
        *a = b as i32;
    }
    sys::bsx_fri_params {
        degree_bits: p.degree_bits as i32,
        rate_bits: p.config.rate_bits as i32,
        cap_height: p.config.cap_height as i32,
        proof_of_work_bits: p.config.proof_of_work_bits as i32,
        num_query_rounds: p.config.num_query_rounds as i32,
        n_arity: p.reduction_arity_bits.len() as i32,
        arity_bits,
    }
}

/// `Challenger<F, PoseidonHash>` <-> the library's plain struct (sponge state, input buffer, output buffer; outputs
/// are popped from the back on both sides).  Needs `pub(crate)` access to the challenger's fields: lives inside the
/// patched plonky2 crate.
pub trait ChallengerBridge<F: RichField> {
    fn to_bsx(&self) -> sys::bsx_challenger;
    fn set_from_bsx(&mut self, c: &sys::bsx_challenger);
}

fn ext<F: RichField + Extendable<D>, const D: usize>(w: &[u64]) -> F::Extension {
    let mut base = [F::ZERO; D];
    for (b, &x) in base.iter_mut().zip(w) {
        *b = F::from_canonical_u64(x);
    }
    F::Extension::from_basefield_array(base)
}

/// The flat stream -> plonky2's `FriProof` (field order and offsets from `bsx_fri_proof_layout`).
pub fn fri_proof_from_words<F: RichField + Extendable<D>, H: Hasher<F>, const D: usize>(
    w: &[u64],
    lay: &sys::bsx_fri_layout,
) -> FriProof<F, H, D>
where
    H::Hash: From<[F; 4]>,
{
    let hashes = |s: &[u64]| -> Vec<H::Hash> { s.chunks(4).map(hash_from_words::<F, H>).collect() };
    let cw = lay.cap_words as usize;
    let commit_phase_merkle_caps = (0..lay.n_rounds as usize)
        .map(|r| MerkleCap(hashes(&w[lay.caps_offset as usize + r * cw..][..cw])))
        .collect();
    let query_round_proofs = (0..lay.n_query_rounds as usize)
        .map(|q| {
            let base = lay.query_offset as usize + q * lay.query_stride as usize;
            let evals_proofs = (0..lay.n_oracles as usize)
                .map(|o| {
                    let leaf = &w[base + lay.initial_leaf_offset[o] as usize..][..lay.initial_leaf_words[o] as usize];
                    let sib = &w[base + lay.initial_proof_offset[o] as usize..][..4 * lay.initial_siblings as usize];
                    (leaf.iter().map(|&x| F::from_canonical_u64(x)).collect(), MerkleProof { siblings: hashes(sib) })
                })
                .collect();
            let steps = (0..lay.n_rounds as usize)
                .map(|r| {
                    let ev = &w[base + lay.step_evals_offset[r] as usize..][..lay.step_evals_words[r] as usize];
                    let sib = &w[base + lay.step_proof_offset[r] as usize..][..4 * lay.step_siblings[r] as usize];
                    FriQueryStep { evals: ev.chunks(2).map(ext::<F, D>).collect(), merkle_proof: MerkleProof { siblings: hashes(sib) } }
                })
                .collect();
            FriQueryRound { initial_trees_proof: FriInitialTreeProof { evals_proofs }, steps }
        })
        .collect();
    let fp = &w[lay.final_poly_offset as usize..][..lay.final_poly_words as usize];
    FriProof {
        commit_phase_merkle_caps,
        query_round_proofs,
        final_poly: PolynomialCoeffs::new(fp.chunks(2).map(ext::<F, D>).collect()),
        pow_witness: F::from_canonical_u64(w[lay.pow_witness_offset as usize]),
    }
}

/// `PolynomialBatch::prove_openings(instance, oracles, challenger, fri_params, timing)`
pub fn prove_openings<F: RichField + Extendable<D>, C: GenericConfig<D, F = F> + 'static, const D: usize>(
    instance: &FriInstanceInfo<F, D>,
    oracles: &[&PolynomialBatch<F, C, D>],
    challenger: &mut Challenger<F, C::Hasher>,
    fri_params: &FriParams,
    _timing: &mut TimingTree,
) -> FriProof<F, C::Hasher, D>
where
    Challenger<F, C::Hasher>: ChallengerBridge<F>,
    <C::Hasher as Hasher<F>>::Hash: From<[F; 4]>,
    C::Hasher: 'static,
{
    assert!(D == 2);
    let params = fri_params(fri_params);
    // FriInstanceInfo.batches -> flat (point, [(oracle, polynomial)]) lists
    let points: Vec<u64> = instance
        .batches
        .iter()
        .flat_map(|b| b.point.to_basefield_array().into_iter().map(|x| x.to_canonical_u64()))
        .collect();
    let n_polys: Vec<i32> = instance.batches.iter().map(|b| b.polynomials.len() as i32).collect();
    let poly_oracle: Vec<i32> = instance.batches.iter().flat_map(|b| b.polynomials.iter().map(|p| p.oracle_index as i32)).collect();
    let poly_index: Vec<i32> = instance.batches.iter().flat_map(|b| b.polynomials.iter().map(|p| p.polynomial_index as i32)).collect();
    let inst = sys::bsx_fri_instance {
        n_batches: instance.batches.len() as i32,
        points: points.as_ptr(),
        n_polys: n_polys.as_ptr(),
        poly_oracle: poly_oracle.as_ptr(),
        poly_index: poly_index.as_ptr(),
    };
    let raws: Vec<*mut sys::bsx_batch> = oracles.iter().map(|o| o.raw.as_ptr()).collect();
    let widths: Vec<i32> = oracles.iter().map(|o| o.merkle_tree.leaf_len as i32).collect();
    let mut lay: sys::bsx_fri_layout = unsafe { std::mem::zeroed() };
    let rc = unsafe { sys::bsx_fri_proof_layout(&params, widths.as_ptr(), widths.len() as i32, &mut lay) };
    assert_eq!(rc, sys::BSX_OK, "FriParams do not describe a valid proof shape");
    let mut words = vec![0u64; lay.total_words as usize];
    let mut ch = challenger.to_bsx();
    with_ctx(|c| {
        c.check(unsafe {
            sys::bsx_fri_prove_openings(c.raw(), raws.as_ptr(), raws.len() as i32, &inst, &mut ch, &params, words.as_mut_ptr())
        })
    });
    challenger.set_from_bsx(&ch); // advanced exactly as upstream's is (alpha, caps, betas, final poly, PoW, queries)
    fri_proof_from_words::<F, C::Hasher, D>(&words, &lay)
}

/// `fri::prover::fri_proof(initial_merkle_trees, lde_polynomial_coeffs, lde_polynomial_values, challenger, fri_params,
/// timing)`.  The values are recomputed on the device from the coefficients (upstream passes both).
pub fn fri_proof<F: RichField + Extendable<D>, C: GenericConfig<D, F = F> + 'static, const D: usize>(
    initial_oracles: &[&PolynomialBatch<F, C, D>],
    lde_polynomial_coeffs: PolynomialCoeffs<F::Extension>,
    challenger: &mut Challenger<F, C::Hasher>,
    fri_params: &FriParams,
    _timing: &mut TimingTree,
) -> FriProof<F, C::Hasher, D>
where
    Challenger<F, C::Hasher>: ChallengerBridge<F>,
    <C::Hasher as Hasher<F>>::Hash: From<[F; 4]>,
    C::Hasher: 'static,
{
    assert!(D == 2);
    let params = fri_params(fri_params);
    assert_eq!(lde_polynomial_coeffs.len(), 1usize << (fri_params.degree_bits + fri_params.config.rate_bits));
    let coeffs: Vec<u64> = lde_polynomial_coeffs
        .coeffs
        .iter()
        .flat_map(|e| e.to_basefield_array().into_iter().map(|x| x.to_canonical_u64()))
        .collect();
    let raws: Vec<*mut sys::bsx_batch> = initial_oracles.iter().map(|o| o.raw.as_ptr()).collect();
    let widths: Vec<i32> = initial_oracles.iter().map(|o| o.merkle_tree.leaf_len as i32).collect();
    let mut lay: sys::bsx_fri_layout = unsafe { std::mem::zeroed() };
    assert_eq!(unsafe { sys::bsx_fri_proof_layout(&params, widths.as_ptr(), widths.len() as i32, &mut lay) }, sys::BSX_OK);
    let mut words = vec![0u64; lay.total_words as usize];
    let mut ch = challenger.to_bsx();
    with_ctx(|c| {
        c.check(unsafe {
            sys::bsx_fri_proof(c.raw(), raws.as_ptr(), raws.len() as i32, coeffs.as_ptr(), &mut ch, &params, words.as_mut_ptr())
        })
    });
    challenger.set_from_bsx(&ch);
    fri_proof_from_words::<F, C::Hasher, D>(&words, &lay)
}
