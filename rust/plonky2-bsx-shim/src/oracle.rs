//! Replaces plonky2/src/fri/oracle.rs: `PolynomialBatch::{from_values, from_coeffs, get_lde_values,
//! get_lde_values_packed}`; `prove_openings` is in prover.rs.
//!
//! Upstream's struct: `{ polynomials: Vec<PolynomialCoeffs<F>>, merkle_tree: MerkleTree<F, C::Hasher>, degree_log,
//! rate_bits, blinding }`.  Here coefficients, LDE leaves and digests stay in HBM behind `raw`; `merkle_tree.cap` is on
//! the host; `polynomials()` / `export()` bring the rest over only when asked (9 GB at 2^20 x 135: PCIe, not HBM, would
//! bound a prover that copied them back after every commit).
use std::ptr::NonNull;

use bsx_b200_sys as sys;
use plonky2::hash::hash_types::RichField;
use plonky2::plonk::config::{GenericConfig, Hasher};
use plonky2::util::timing::TimingTree;
use plonky2_field::extension::Extendable;
use plonky2_field::fft::FftRootTable;
use plonky2_field::polynomial::{PolynomialCoeffs, PolynomialValues};
use plonky2_util::log2_strict;

use crate::merkle_tree::DeviceMerkleTree;
use crate::{assert_goldilocks_poseidon, with_ctx, words_to_field};

pub struct PolynomialBatch<F: RichField + Extendable<D>, C: GenericConfig<D, F = F>, const D: usize> {
    pub(crate) raw: NonNull<sys::bsx_batch>,
    pub merkle_tree: DeviceMerkleTree<F, C::Hasher>,
    pub degree_log: usize,
    pub rate_bits: usize,
    pub blinding: bool,
    pub(crate) n_polys: usize,
}

impl<F: RichField + Extendable<D>, C: GenericConfig<D, F = F> + 'static, const D: usize> PolynomialBatch<F, C, D>
where
    <C::Hasher as Hasher<F>>::Hash: From<[F; 4]>,
    C::Hasher: 'static,
{
    /// `PolynomialBatch::from_values(values, rate_bits, blinding, cap_height, timing, fft_root_table)`.
    /// `fft_root_table` is ignored: the context caches its own twiddle and coset tables per shape.
    pub fn from_values(
        values: Vec<PolynomialValues<F>>,
        rate_bits: usize,
        blinding: bool,
        cap_height: usize,
        timing: &mut TimingTree,
        _fft_root_table: Option<&FftRootTable<F>>,
    ) -> Self {
        Self::build(values.iter().map(|p| p.values.as_slice()).collect(), rate_bits, blinding, cap_height, timing, false)
    }

    /// `PolynomialBatch::from_coeffs(polynomials, rate_bits, blinding, cap_height, timing, fft_root_table)`
    pub fn from_coeffs(
        polynomials: Vec<PolynomialCoeffs<F>>,
        rate_bits: usize,
        blinding: bool,
        cap_height: usize,
        timing: &mut TimingTree,
        _fft_root_table: Option<&FftRootTable<F>>,
    ) -> Self {
        Self::build(polynomials.iter().map(|p| p.coeffs.as_slice()).collect(), rate_bits, blinding, cap_height, timing, true)
    }

    fn build(cols: Vec<&[F]>, rate_bits: usize, blinding: bool, cap_height: usize, timing: &mut TimingTree, coeffs: bool) -> Self {
        assert_goldilocks_poseidon::<F, C::Hasher>();
        assert_eq!(D, 2, "bsx_b200 implements the quadratic extension only");
        let n = cols.first().map_or(1, |c| c.len());
        let degree_log = log2_strict(n); // upstream panics on a non power of two here as well
        assert!(cols.iter().all(|c| c.len() == n), "polynomials of different lengths");
        // No gather: every column is handed over where it lies in the heap (one pointer per polynomial, pageable
        // memory); the library stages the columns through its pinned ring while earlier chunks are being transformed.
        let ptrs: Vec<*const u64> = cols.iter().map(|c| c.as_ptr() as *const u64).collect();
        let mut raw = std::ptr::null_mut();
        let mut t = sys::bsx_timing { h2d_ms: 0.0, ifft_ms: 0.0, lde_ms: 0.0, leaf_hash_ms: 0.0, tree_ms: 0.0, total_ms: 0.0 };
        with_ctx(|c| {
            let rc = unsafe {
                if coeffs {
                    sys::bsx_batch_from_coeffs_cols(c.raw(), ptrs.as_ptr(), degree_log as i32, cols.len() as i32, rate_bits as i32,
                                                    cap_height as i32, blinding as i32, std::ptr::null(), &mut raw, &mut t)
                } else {
                    sys::bsx_batch_from_values_cols(c.raw(), ptrs.as_ptr(), degree_log as i32, cols.len() as i32, rate_bits as i32,
                                                    cap_height as i32, blinding as i32, std::ptr::null(), &mut raw, &mut t)
                }
            };
            c.check(rc)
        });
        // the scopes upstream opens with timed!(): "IFFT", "FFT + blinding", "build Merkle tree"
        timing.push("IFFT", log::Level::Debug);
        timing.pop();
        let _ = (t.ifft_ms, t.lde_ms, t.leaf_hash_ms + t.tree_ms); // device milliseconds; a fork of TimingTree can take them verbatim
        Self::from_raw(NonNull::new(raw).unwrap(), blinding)
    }

    pub(crate) fn from_raw(raw: NonNull<sys::bsx_batch>, blinding: bool) -> Self {
        let (mut log_n, mut n_cols, mut rate_bits, mut cap_height, mut n_salt) = (0i32, 0i32, 0i32, 0i32, 0i32);
        unsafe { sys::bsx_batch_shape(raw.as_ptr(), &mut log_n, &mut n_cols, &mut rate_bits, &mut cap_height, &mut n_salt) };
        let tree = NonNull::new(unsafe { sys::bsx_batch_tree(raw.as_ptr()) }).unwrap();
        Self {
            raw,
            merkle_tree: DeviceMerkleTree::from_raw(tree, false),
            degree_log: log_n as usize,
            rate_bits: rate_bits as usize,
            blinding,
            n_polys: n_cols as usize,
        }
    }

    /// `PolynomialBatch::get_lde_values(index, step)`
    pub fn get_lde_values(&self, index: usize, step: usize) -> Vec<F> {
        let mut out = vec![0u64; self.n_polys];
        with_ctx(|c| c.check(unsafe { sys::bsx_batch_get_lde_values(self.raw.as_ptr(), index as u64, step as u64, out.as_mut_ptr()) }));
        words_to_field(out)
    }

    /// `PolynomialBatch::get_lde_values_packed(index_start, step)` for `count` consecutive indices: the rows, one after
    /// the other (upstream transposes them into P::WIDTH-lane packed columns; the CPU quotient evaluator is the only caller).
    pub fn get_lde_values_rows(&self, index_start: usize, count: usize, step: usize) -> Vec<F> {
        let mut out = vec![0u64; self.n_polys * count];
        with_ctx(|c| {
            c.check(unsafe {
                sys::bsx_batch_get_lde_values_packed(self.raw.as_ptr(), index_start as u64, count as u64, step as u64,
                                                     out.as_mut_ptr(), sys::BSX_HOST)
            })
        });
        words_to_field(out)
    }

    /// upstream's public field `polynomials`, materialised on request (natural order)
    pub fn polynomials(&self) -> Vec<PolynomialCoeffs<F>> {
        let n = 1usize << self.degree_log;
        let mut w = vec![0u64; n * self.n_polys];
        with_ctx(|c| c.check(unsafe { sys::bsx_batch_coeffs(self.raw.as_ptr(), w.as_mut_ptr(), sys::BSX_HOST) }));
        let all: Vec<F> = words_to_field(w);
        all.chunks(n).map(|c| PolynomialCoeffs::new(c.to_vec())).collect()
    }

    /// `eval_commitment(z, c)` of plonk/proof.rs `OpeningSet::new`: every polynomial at extension points, on the device
    pub fn eval_commitment(&self, points: &[[u64; 2]]) -> Vec<Vec<[u64; 2]>> {
        let mut out = vec![0u64; 2 * self.n_polys * points.len()];
        with_ctx(|c| {
            c.check(unsafe {
                sys::bsx_batch_eval_ext(self.raw.as_ptr(), 0, self.n_polys as i32, points.as_ptr() as *const u64,
                                        points.len() as i32, out.as_mut_ptr())
            })
        });
        out.chunks(2 * self.n_polys).map(|p| p.chunks(2).map(|e| [e[0], e[1]]).collect()).collect()
    }
}

impl<F: RichField + Extendable<D>, C: GenericConfig<D, F = F>, const D: usize> Drop for PolynomialBatch<F, C, D> {
    fn drop(&mut self) {
        unsafe { sys::bsx_batch_free(self.raw.as_ptr()) }
    }
}
