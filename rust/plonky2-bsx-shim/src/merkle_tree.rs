//! Replaces plonky2/src/hash/merkle_tree.rs: `MerkleTree::new`, `get`, `prove`.
//!
//! Upstream's struct has public `leaves: Vec<Vec<F>>`, `digests: Vec<H::Hash>`, `cap: MerkleCap<F, H>`.  The patched
//! struct keeps `cap` on the host and serves leaves, digests and proofs from the device-resident tree on demand
//! (`leaves()` / `digests()` materialise them only when a caller -- the circuit serialiser -- really wants them).
use std::marker::PhantomData;
use std::ptr::NonNull;

use bsx_b200_sys as sys;
use plonky2::hash::hash_types::RichField;
use plonky2::hash::merkle_proofs::MerkleProof;
use plonky2::hash::merkle_tree::MerkleCap;
use plonky2::plonk::config::Hasher;
use plonky2_util::log2_strict;

use crate::{as_words, assert_goldilocks_poseidon, hash_from_words, with_ctx, words_to_field};

pub struct DeviceMerkleTree<F: RichField, H: Hasher<F>> {
    pub(crate) raw: NonNull<sys::bsx_tree>,
    /// false for the tree embedded in a PolynomialBatch (freed with the batch)
    pub(crate) owned: bool,
    pub cap: MerkleCap<F, H>,
    pub(crate) log_leaves: usize,
    pub(crate) leaf_len: usize,
    pub(crate) cap_height: usize,
    _p: PhantomData<H>,
}

impl<F: RichField, H: Hasher<F> + 'static> DeviceMerkleTree<F, H>
where
    H::Hash: From<[F; 4]>,
{
    /// `MerkleTree::new(leaves, cap_height)`
    pub fn new(leaves: Vec<Vec<F>>, cap_height: usize) -> Self {
        assert_goldilocks_poseidon::<F, H>();
        let log_leaves = log2_strict(leaves.len()); // panics like upstream on a non power of two
        let leaf_len = leaves.first().map_or(0, |l| l.len());
        assert!(leaves.iter().all(|l| l.len() == leaf_len), "leaves of different lengths");
        // row-major staging copy: upstream's Vec<Vec<F>> is one heap allocation per leaf
        let mut flat: Vec<F> = Vec::with_capacity(leaves.len() * leaf_len);
        for l in &leaves {
            flat.extend_from_slice(l);
        }
        let mut raw = std::ptr::null_mut();
        with_ctx(|c| {
            // the library repeats upstream's assert text: "cap_height=.. should be at most log2(leaves.len())=.."
            c.check(unsafe {
                sys::bsx_merkle_new(c.raw(), as_words(&flat), sys::BSX_HOST, log_leaves as i32, leaf_len as i32,
                                    cap_height as i32, &mut raw)
            })
        });
        Self::from_raw(NonNull::new(raw).unwrap(), true)
    }

    pub(crate) fn from_raw(raw: NonNull<sys::bsx_tree>, owned: bool) -> Self {
        let (mut ll, mut len, mut ch) = (0i32, 0i32, 0i32);
        unsafe { sys::bsx_merkle_shape(raw.as_ptr(), &mut ll, &mut len, &mut ch) };
        let mut cap_words = vec![0u64; 4usize << ch];
        unsafe { sys::bsx_merkle_cap(raw.as_ptr(), cap_words.as_mut_ptr()) };
        let cap = MerkleCap(cap_words.chunks(4).map(hash_from_words::<F, H>).collect());
        Self { raw, owned, cap, log_leaves: ll as usize, leaf_len: len as usize, cap_height: ch as usize, _p: PhantomData }
    }

    /// `MerkleTree::get(i)`
    pub fn get(&self, i: usize) -> Vec<F> {
        let mut out = vec![0u64; self.leaf_len];
        let idx = [i as u64];
        with_ctx(|c| c.check(unsafe { sys::bsx_merkle_get(self.raw.as_ptr(), idx.as_ptr(), 1, out.as_mut_ptr()) }));
        words_to_field(out)
    }

    /// `MerkleTree::prove(leaf_index)`
    pub fn prove(&self, leaf_index: usize) -> MerkleProof<F, H> {
        let n_sib = self.log_leaves - self.cap_height;
        let mut sib = vec![0u64; 4 * n_sib];
        with_ctx(|c| c.check(unsafe { sys::bsx_merkle_prove(self.raw.as_ptr(), leaf_index as u64, sib.as_mut_ptr()) }));
        MerkleProof { siblings: sib.chunks(4).map(hash_from_words::<F, H>).collect() }
    }

    /// `MerkleTree.digests` in upstream's recursive layout (only the circuit serialiser reads it)
    pub fn digests(&self) -> Vec<H::Hash> {
        let n = 2 * ((1usize << self.log_leaves) - (1usize << self.cap_height));
        let mut w = vec![0u64; 4 * n];
        with_ctx(|c| c.check(unsafe { sys::bsx_merkle_digests(self.raw.as_ptr(), w.as_mut_ptr()) }));
        w.chunks(4).map(hash_from_words::<F, H>).collect()
    }
}

impl<F: RichField, H: Hasher<F>> Drop for DeviceMerkleTree<F, H> {
    fn drop(&mut self) {
        if self.owned {
            unsafe { sys::bsx_merkle_free(self.raw.as_ptr()) }
        }
    }
}
