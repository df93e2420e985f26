"""Host mirror of plonky2 plonk/proof.rs: `OpeningSet`, `Proof`, `ProofWithPublicInputs` and their byte form
(`ProofWithPublicInputs::to_bytes` / `from_bytes`, util/serialization/mod.rs `write_proof_with_public_inputs`), over the
C ABI entry points bsx_fri_proof_layout / bsx_fri_proof_bytes / bsx_proof_with_public_inputs_bytes.  F = Goldilocks,
D = 2, Poseidon.  The byte layout is restated from the public crate (unpinned: the reference mount holds no bytes)."""
import ctypes as C
from dataclasses import dataclass, field
from typing import List

import numpy as np

from .binding import BsxError, lib, _np, _ptr
from .fri import _params_struct

BSX_MAX_ORACLES = 16
OPENING_SET_FIELDS = ("constants", "plonk_sigmas", "wires", "plonk_zs", "plonk_zs_next", "partial_products",
                      "quotient_polys")


class FriLayout(C.Structure):
    """include/bsx_b200.h `bsx_fri_layout`."""
    _fields_ = [("total_words", C.c_uint64), ("caps_offset", C.c_uint64), ("cap_words", C.c_uint64),
                ("query_offset", C.c_uint64), ("query_stride", C.c_uint64),
                ("initial_leaf_offset", C.c_uint64 * BSX_MAX_ORACLES), ("initial_leaf_words", C.c_uint64 * BSX_MAX_ORACLES),
                ("initial_proof_offset", C.c_uint64 * BSX_MAX_ORACLES),
                ("step_evals_offset", C.c_uint64 * 16), ("step_evals_words", C.c_uint64 * 16),
                ("step_proof_offset", C.c_uint64 * 16),
                ("final_poly_offset", C.c_uint64), ("final_poly_words", C.c_uint64), ("pow_witness_offset", C.c_uint64),
                ("n_oracles", C.c_int32), ("n_rounds", C.c_int32), ("n_query_rounds", C.c_int32),
                ("initial_siblings", C.c_int32), ("step_siblings", C.c_int32 * 16)]


class _ProofParts(C.Structure):
    _fields_ = [("wires_cap", C.c_void_p), ("plonk_zs_partial_products_cap", C.c_void_p), ("quotient_polys_cap", C.c_void_p),
                ("openings", C.c_void_p * 7), ("n_openings", C.c_int32 * 7), ("opening_proof_words", C.c_void_p),
                ("public_inputs", C.c_void_p), ("n_public_inputs", C.c_int32)]


def _declare():
    L = lib()
    if getattr(L, "_bsx_proof_declared", False):
        return L
    vp, i32, sz = C.c_void_p, C.c_int32, C.c_size_t
    L.bsx_fri_proof_layout.restype = i32
    L.bsx_fri_proof_layout.argtypes = [vp, vp, i32, C.POINTER(FriLayout)]
    L.bsx_fri_proof_bytes.restype = i32
    L.bsx_fri_proof_bytes.argtypes = [vp, vp, vp, i32, vp, sz, C.POINTER(sz)]
    L.bsx_proof_with_public_inputs_bytes.restype = i32
    L.bsx_proof_with_public_inputs_bytes.argtypes = [C.POINTER(_ProofParts), vp, vp, i32, vp, sz, C.POINTER(sz)]
    L._bsx_proof_declared = True
    return L


def _check(rc, what):
    if rc != 0:
        raise BsxError("%s failed (code %d)" % (what, rc))


def fri_proof_layout(fri_params, leaf_widths):
    """Word offsets of every `FriProof` field in the flat stream (bsx_fri_proof_layout)."""
    w = np.ascontiguousarray(leaf_widths, dtype=np.int32)
    out = FriLayout()
    p = _params_struct(fri_params)
    _check(_declare().bsx_fri_proof_layout(C.byref(p), C.c_void_p(w.ctypes.data), int(w.size), C.byref(out)),
           "bsx_fri_proof_layout")
    return out


def fri_proof_bytes(proof_words, fri_params, leaf_widths):
    """`Write::write_fri_proof` through the C ABI."""
    w = np.ascontiguousarray(leaf_widths, dtype=np.int32)
    words = _np(proof_words).reshape(-1)
    p = _params_struct(fri_params)
    n = C.c_size_t()
    L = _declare()
    _check(L.bsx_fri_proof_bytes(_ptr(words), C.byref(p), C.c_void_p(w.ctypes.data), int(w.size), None, 0, C.byref(n)),
           "bsx_fri_proof_bytes")
    buf = np.zeros(n.value, dtype=np.uint8)
    _check(L.bsx_fri_proof_bytes(_ptr(words), C.byref(p), C.c_void_p(w.ctypes.data), int(w.size),
                                 C.c_void_p(buf.ctypes.data), buf.size, C.byref(n)), "bsx_fri_proof_bytes")
    return buf.tobytes()


@dataclass
class OpeningSet:
    """plonk/proof.rs `OpeningSet<F, D>`: every field an (n, 2) array of extension elements."""
    constants: np.ndarray
    plonk_sigmas: np.ndarray
    wires: np.ndarray
    plonk_zs: np.ndarray
    plonk_zs_next: np.ndarray
    partial_products: np.ndarray
    quotient_polys: np.ndarray

    @classmethod
    def new(cls, zeta, g, constants_sigmas_commitment, wires_commitment, zs_partial_products_commitment,
            quotient_polys_commitment, num_constants, num_challenges):
        """`OpeningSet::new(zeta, g, ..commitments.., common_data)`: `eval_commitment` at zeta (and the Zs at g zeta).
        num_constants = common_data.num_constants (selectors included): constants_range = 0..num_constants, sigmas
        after it; zs_range = 0..num_challenges, partial products after it."""
        P = 0xFFFFFFFF00000001
        zeta = _np(zeta).reshape(2)
        gz = np.array([int(zeta[0]) * int(g) % P, int(zeta[1]) * int(g) % P], dtype=np.uint64)
        cs = constants_sigmas_commitment.eval_commitment(zeta)
        zs = zs_partial_products_commitment.eval_commitment(np.stack([zeta, gz]), 0, None)
        return cls(constants=cs[:num_constants], plonk_sigmas=cs[num_constants:],
                   wires=wires_commitment.eval_commitment(zeta), plonk_zs=zs[0][:num_challenges],
                   plonk_zs_next=zs[1][:num_challenges], partial_products=zs[0][num_challenges:],
                   quotient_polys=quotient_polys_commitment.eval_commitment(zeta))

    def to_fri_openings(self):
        """`OpeningSet::to_fri_openings`: [zeta batch, zeta_next batch] in the order the challenger observes them."""
        zeta = np.concatenate([self.constants, self.plonk_sigmas, self.wires, self.plonk_zs, self.partial_products,
                               self.quotient_polys]).reshape(-1, 2)
        return [zeta, np.asarray(self.plonk_zs_next).reshape(-1, 2)]

    def fields(self):
        return [np.ascontiguousarray(getattr(self, k), dtype=np.uint64).reshape(-1, 2) for k in OPENING_SET_FIELDS]


@dataclass
class Proof:
    """plonk/proof.rs `Proof<F, C, D>`; opening_proof is the flat FriProof stream of prove_openings."""
    wires_cap: np.ndarray
    plonk_zs_partial_products_cap: np.ndarray
    quotient_polys_cap: np.ndarray
    openings: OpeningSet
    opening_proof: np.ndarray


@dataclass
class ProofWithPublicInputs:
    """plonk/proof.rs `ProofWithPublicInputs<F, C, D>`."""
    proof: Proof
    public_inputs: np.ndarray = field(default_factory=lambda: np.zeros(0, dtype=np.uint64))

    def to_bytes(self, fri_params, leaf_widths):
        """`ProofWithPublicInputs::to_bytes()`.  leaf_widths: polynomials + salt columns of the four oracles
        (constants_sigmas, wires, zs_partial_products, quotient) in the order prove_openings got them."""
        pr = self.proof
        keep = [_np(pr.wires_cap), _np(pr.plonk_zs_partial_products_cap), _np(pr.quotient_polys_cap)] + pr.openings.fields() + \
               [_np(pr.opening_proof).reshape(-1), _np(self.public_inputs).reshape(-1)]
        parts = _ProofParts()
        parts.wires_cap, parts.plonk_zs_partial_products_cap, parts.quotient_polys_cap = (a.ctypes.data for a in keep[:3])
        for k in range(7):
            parts.openings[k] = keep[3 + k].ctypes.data
            parts.n_openings[k] = keep[3 + k].shape[0]
        parts.opening_proof_words = keep[10].ctypes.data
        parts.public_inputs = keep[11].ctypes.data if keep[11].size else None
        parts.n_public_inputs = int(keep[11].size)
        w = np.ascontiguousarray(leaf_widths, dtype=np.int32)
        lay = fri_proof_layout(fri_params, w)
        if keep[10].size != lay.total_words:
            raise BsxError("opening proof has %d words, expected %d" % (keep[10].size, lay.total_words))
        for a in keep[:3]:
            if a.size != lay.cap_words:
                raise BsxError("a cap has %d words, expected %d" % (a.size, lay.cap_words))
        p = _params_struct(fri_params)
        n = C.c_size_t()
        L = _declare()
        _check(L.bsx_proof_with_public_inputs_bytes(C.byref(parts), C.byref(p), C.c_void_p(w.ctypes.data), int(w.size),
                                                    None, 0, C.byref(n)), "bsx_proof_with_public_inputs_bytes")
        buf = np.zeros(n.value, dtype=np.uint8)
        _check(L.bsx_proof_with_public_inputs_bytes(C.byref(parts), C.byref(p), C.c_void_p(w.ctypes.data), int(w.size),
                                                    C.c_void_p(buf.ctypes.data), buf.size, C.byref(n)),
               "bsx_proof_with_public_inputs_bytes")
        return buf.tobytes()

    @classmethod
    def from_bytes(cls, data, fri_params, leaf_widths, n_openings):
        """`ProofWithPublicInputs::from_bytes(bytes, common_data)`: n_openings = the seven `OpeningSet` vector lengths
        (upstream takes them from CommonCircuitData); the public inputs are whatever follows the proof."""
        from .serialization import read_fri_proof
        buf = memoryview(data).cast("B")
        c = fri_params.config
        cap_bytes = 32 << c.cap_height
        pos = 0

        def take(nbytes):
            nonlocal pos
            if pos + nbytes > len(buf):
                raise ValueError("proof bytes truncated")
            a = np.frombuffer(buf[pos:pos + nbytes], dtype="<u8").astype(np.uint64)
            pos += nbytes
            return a
        caps = [take(cap_bytes).reshape(-1, 4) for _ in range(3)]
        ops = [take(16 * int(k)).reshape(-1, 2) for k in n_openings]
        lay = fri_proof_layout(fri_params, leaf_widths)
        fri_len = 8 * lay.total_words + c.num_query_rounds * (lay.n_oracles + lay.n_rounds)  # one length byte per Merkle proof
        if pos + fri_len > len(buf):
            raise ValueError("proof bytes truncated")
        words = read_fri_proof(bytes(buf[pos:pos + fri_len]), list(leaf_widths), fri_params.degree_bits, c.rate_bits,
                               c.cap_height, list(fri_params.reduction_arity_bits), c.num_query_rounds)
        pos += fri_len
        if (len(buf) - pos) % 8:
            raise ValueError("public inputs are not whole field elements")
        pis = take(len(buf) - pos)
        return cls(Proof(caps[0], caps[1], caps[2], OpeningSet(*ops), words), pis)
