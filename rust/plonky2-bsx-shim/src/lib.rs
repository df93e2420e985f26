//! plonky2-bsx-shim: the bodies a patched plonky2 delegates to (Cargo `[patch]`, see rust/README.md).
//!
//! plonky2 has no FFI for its commitment phase: `plonk/prover.rs::prove` calls `PolynomialBatch::from_values`,
//! `MerkleTree::new`, `PolynomialBatch::prove_openings`, `fri_proof` directly.  A drop-in keeps those signatures
//! and swaps their bodies for the calls below.  Everything here is specialised to
//! `F = GoldilocksField`, `C::Hasher = PoseidonHash`, `D = 2` and refuses anything else by panicking --
//! there is no CPU fallback, exactly like the library (`bsx_ctx_create` fails without a CUDA device).
//!
//! Modules mirror the upstream files they patch:
//!   oracle.rs       -> plonky2/src/fri/oracle.rs        (`PolynomialBatch`)
//!   merkle_tree.rs  -> plonky2/src/hash/merkle_tree.rs  (`MerkleTree`)
//!   prover.rs       -> plonky2/src/fri/prover.rs        (`fri_proof`), fri/oracle.rs `prove_openings`
//!   quotient.rs     -> plonky2/src/plonk/prover.rs      (`compute_quotient_polys`, partial products and Zs)
//!
//! This image has no rustc / cargo, so the crate is unbuilt here; the same entry points are exercised by the ctypes
//! mirror (0-kno-blobstreamx_b200/*.py) and the plain-C harness (tests/c_harness), and tests/test_abi.py checks that
//! every `bsx_*` call made from these files exists in the header with the same number of arguments.

pub mod merkle_tree;
pub mod oracle;
pub mod prover;
pub mod quotient;

use std::any::TypeId;
use std::cell::RefCell;
use std::ffi::CStr;

use bsx_b200_sys as sys;
use plonky2::hash::poseidon::PoseidonHash;
use plonky2::plonk::config::Hasher;
use plonky2_field::goldilocks_field::GoldilocksField;
use plonky2_field::types::Field;

/// One `bsx_ctx` per (host thread, device): plonky2x map-reduce runs several proofs in one process, each on its own
/// thread; contexts are independent and the library has no global mutable state.
pub struct Context {
    raw: *mut sys::bsx_ctx,
}

impl Context {
    pub fn new(device: i32) -> Self {
        let mut raw = std::ptr::null_mut();
        let rc = unsafe { sys::bsx_ctx_create(device, &mut raw) };
        if rc != sys::BSX_OK {
            panic!("bsx_ctx_create(device = {device}) failed with {rc}: no usable CUDA device (there is no CPU fallback)");
        }
        Self { raw }
    }
    pub fn raw(&self) -> *mut sys::bsx_ctx {
        self.raw
    }
    /// Upstream's functions are infallible by signature and panic on misuse; a non-zero status becomes that panic,
    /// with the library's message (which repeats upstream's text where upstream has one).
    pub fn check(&self, rc: i32) {
        if rc != sys::BSX_OK {
            let msg = unsafe { CStr::from_ptr(sys::bsx_last_error(self.raw)) }.to_string_lossy().into_owned();
            panic!("{msg} (bsx status {rc})");
        }
    }
}

impl Drop for Context {
    fn drop(&mut self) {
        unsafe { sys::bsx_ctx_destroy(self.raw) }
    }
}

thread_local! {
    static CTX: RefCell<Option<Context>> = const { RefCell::new(None) };
}

/// The calling thread's context on the device named by `BSX_B200_DEVICE` (default 0), created on first use.
pub fn with_ctx<R>(f: impl FnOnce(&Context) -> R) -> R {
    CTX.with(|c| {
        let mut c = c.borrow_mut();
        if c.is_none() {
            let dev = std::env::var("BSX_B200_DEVICE").ok().and_then(|s| s.parse().ok()).unwrap_or(0);
            *c = Some(Context::new(dev));
        }
        f(c.as_ref().unwrap())
    })
}

/// The only instantiation the library implements.  Anything else is refused, not emulated.
pub fn assert_goldilocks_poseidon<F: 'static, H: 'static>() {
    assert!(
        TypeId::of::<F>() == TypeId::of::<GoldilocksField>() && TypeId::of::<H>() == TypeId::of::<PoseidonHash>(),
        "bsx_b200 implements F = GoldilocksField with PoseidonHash only; this instantiation has no device path and no fallback"
    );
}

/// `GoldilocksField` is `#[repr(transparent)]` over `u64`; the library accepts any u64 representative and returns
/// canonical ones, which is what `GoldilocksField(u64)` holds after `to_canonical`.
pub fn as_words<F: Field + 'static>(v: &[F]) -> *const u64 {
    debug_assert!(TypeId::of::<F>() == TypeId::of::<GoldilocksField>());
    v.as_ptr() as *const u64
}
pub fn words_to_field<F: Field + 'static>(w: Vec<u64>) -> Vec<F> {
    debug_assert!(TypeId::of::<F>() == TypeId::of::<GoldilocksField>());
    let mut w = std::mem::ManuallyDrop::new(w);
    unsafe { Vec::from_raw_parts(w.as_mut_ptr() as *mut F, w.len(), w.capacity()) }
}
pub fn hash_from_words<F: Field + 'static, H: Hasher<F>>(w: &[u64]) -> H::Hash
where
    H::Hash: From<[F; 4]>,
{
    let e: Vec<F> = w[..4].iter().map(|&x| F::from_canonical_u64(x)).collect();
    H::Hash::from([e[0], e[1], e[2], e[3]])
}
