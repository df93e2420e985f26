"""Host-side FRI configuration mirror (plonky2 fri/mod.rs, fri/reduction_strategies.rs)."""
from dataclasses import dataclass, field
from typing import List


def fri_reduction_arity_bits(degree_bits, rate_bits, cap_height, arity_bits=4, final_poly_bits=5):
    """`FriReductionStrategy::ConstantArityBits(arity_bits, final_poly_bits).reduction_arity_bits(..)`."""
    out = []
    while degree_bits > final_poly_bits and degree_bits + rate_bits - arity_bits >= cap_height:
        out.append(arity_bits)
        assert degree_bits >= arity_bits
        degree_bits -= arity_bits
    return out


@dataclass
class FriConfig:
    """plonky2 `FriConfig`; defaults are `CircuitConfig::standard_recursion_config().fri_config`."""
    rate_bits: int = 3
    cap_height: int = 4
    proof_of_work_bits: int = 16
    arity_bits: int = 4
    final_poly_bits: int = 5
    num_query_rounds: int = 28

    def fri_params(self, degree_bits, hiding=False):
        return FriParams(self, hiding, degree_bits,
                         fri_reduction_arity_bits(degree_bits, self.rate_bits, self.cap_height, self.arity_bits,
                                                  self.final_poly_bits))


@dataclass
class FriParams:
    """plonky2 `FriParams`."""
    config: FriConfig
    hiding: bool
    degree_bits: int
    reduction_arity_bits: List[int] = field(default_factory=list)

    def total_arities(self):
        return sum(self.reduction_arity_bits)

    def lde_bits(self):
        return self.degree_bits + self.config.rate_bits

    def final_poly_bits(self):
        return self.degree_bits - self.total_arities()

    def final_poly_len(self):
        return 1 << self.final_poly_bits()


# ---------------------------------------------------------------------------------------------
# Challenger and FRI prover over the C ABI (plonky2 iop/challenger.rs, fri/oracle.rs, fri/prover.rs)
import ctypes as C

import numpy as np

from .binding import BsxError, lib, _np, _ptr


class _ChallengerStruct(C.Structure):
    _fields_ = [("sponge_state", C.c_uint64 * 12), ("input_buffer", C.c_uint64 * 8),
                ("output_buffer", C.c_uint64 * 8), ("n_input", C.c_int32), ("n_output", C.c_int32)]


class Challenger:
    """plonky2 `Challenger<GoldilocksField, PoseidonHash>`; permutations run on the device."""

    def __init__(self, ctx):
        self.ctx = ctx
        self._s = _ChallengerStruct()
        lib().bsx_challenger_init(C.byref(self._s))

    def observe_elements(self, elems):
        e = _np(np.atleast_1d(elems)).reshape(-1)
        self.ctx.check(lib().bsx_challenger_observe(self.ctx._h, C.byref(self._s), _ptr(e), e.size))

    observe_element = observe_elements
    observe_hash = observe_elements

    def observe_cap(self, cap):
        self.observe_elements(cap)

    def observe_extension_elements(self, elems):
        self.observe_elements(elems)

    def get_n_challenges(self, n):
        out = np.zeros(n, dtype=np.uint64)
        self.ctx.check(lib().bsx_challenger_get(self.ctx._h, C.byref(self._s), _ptr(out), n))
        return out

    def get_challenge(self):
        return int(self.get_n_challenges(1)[0])

    def get_extension_challenge(self):
        return self.get_n_challenges(2)

    def copy(self):
        c = Challenger(self.ctx)
        C.memmove(C.byref(c._s), C.byref(self._s), C.sizeof(_ChallengerStruct))
        return c

    def snapshot(self):
        s = self._s
        return (list(s.sponge_state), list(s.input_buffer)[:s.n_input], list(s.output_buffer)[:s.n_output])


class _InstanceStruct(C.Structure):
    _fields_ = [("n_batches", C.c_int32), ("points", C.c_void_p), ("n_polys", C.c_void_p),
                ("poly_oracle", C.c_void_p), ("poly_index", C.c_void_p)]


class _ParamsStruct(C.Structure):
    _fields_ = [("degree_bits", C.c_int32), ("rate_bits", C.c_int32), ("cap_height", C.c_int32),
                ("proof_of_work_bits", C.c_int32), ("num_query_rounds", C.c_int32), ("n_arity", C.c_int32),
                ("arity_bits", C.c_int32 * 16)]


class FriInstanceInfo:
    """plonky2 `FriInstanceInfo`: batches of (opening point, [(oracle_index, polynomial_index)])."""

    def __init__(self, points, polys):
        self.points = _np(points).reshape(-1, 2)
        self.polys = [list(p) for p in polys]
        if len(self.polys) != self.points.shape[0]:
            raise BsxError("one point per opening batch")
        self._n = np.array([len(p) for p in self.polys], dtype=np.int32)
        self._o = np.array([o for p in self.polys for o, _ in p], dtype=np.int32)
        self._i = np.array([i for p in self.polys for _, i in p], dtype=np.int32)
        self.struct = _InstanceStruct(len(self.polys), self.points.ctypes.data, self._n.ctypes.data,
                                      self._o.ctypes.data, self._i.ctypes.data)


def _params_struct(fp: "FriParams"):
    p = _ParamsStruct(fp.degree_bits, fp.config.rate_bits, fp.config.cap_height, fp.config.proof_of_work_bits,
                      fp.config.num_query_rounds, len(fp.reduction_arity_bits))
    for i, a in enumerate(fp.reduction_arity_bits):
        p.arity_bits[i] = a
    return p


def _handles(oracles):
    arr = (C.c_void_p * len(oracles))(*[o._h for o in oracles])
    return arr


def fri_proof_words(oracles, fri_params):
    n = C.c_size_t()
    p = _params_struct(fri_params)
    ctx = oracles[0].ctx
    ctx.check(lib().bsx_fri_proof_words(_handles(oracles), len(oracles), C.byref(p), C.byref(n)))
    return n.value


def prove_openings(instance, oracles, challenger, fri_params):
    """`PolynomialBatch::prove_openings(instance, oracles, challenger, fri_params, timing)` -> flat FriProof."""
    ctx = oracles[0].ctx
    p = _params_struct(fri_params)
    proof = np.zeros(fri_proof_words(oracles, fri_params), dtype=np.uint64)
    ctx.check(lib().bsx_fri_prove_openings(ctx._h, _handles(oracles), len(oracles), C.byref(instance.struct),
                                           C.byref(challenger._s), C.byref(p), _ptr(proof)))
    return proof


def fri_proof_struct(proof_words, oracles, fri_params):
    """The flat proof as plonky2's `FriProof` fields (serialization.fri_proof_struct with the oracles' shapes)."""
    from .serialization import fri_proof_struct as cut
    widths = [b.shape()["n_cols"] + b.shape()["n_salt"] for b in oracles]
    c = fri_params.config
    return cut(proof_words, widths, fri_params.degree_bits, c.rate_bits, c.cap_height,
               list(fri_params.reduction_arity_bits), c.num_query_rounds)


def fri_proof_to_bytes(proof_words, oracles, fri_params):
    """`Write::write_fri_proof` bytes of a proof returned by prove_openings / fri_proof."""
    from .serialization import write_fri_proof
    return write_fri_proof(fri_proof_struct(proof_words, oracles, fri_params))


def fri_proof(oracles, lde_polynomial_coeffs, challenger, fri_params):
    """`fri::prover::fri_proof(initial_merkle_trees, lde_polynomial_coeffs, lde_polynomial_values, ..)`;
    coefficients: (2^(degree_bits + rate_bits), 2) extension elements.  The values are recomputed on device."""
    ctx = oracles[0].ctx
    p = _params_struct(fri_params)
    co = _np(lde_polynomial_coeffs)
    if co.shape != (1 << fri_params.lde_bits(), 2):
        raise BsxError("lde_polynomial_coeffs must have 2^(degree_bits + rate_bits) extension coefficients")
    proof = np.zeros(fri_proof_words(oracles, fri_params), dtype=np.uint64)
    ctx.check(lib().bsx_fri_proof(ctx._h, _handles(oracles), len(oracles), _ptr(co), C.byref(challenger._s),
                                  C.byref(p), _ptr(proof)))
    return proof


def final_poly(instance, oracles, alpha):
    ctx = oracles[0].ctx
    n = 1 << oracles[0].shape()["degree_log"]
    out = np.zeros((n, 2), dtype=np.uint64)
    a = _np(alpha)
    ctx.check(lib().bsx_fri_final_poly(ctx._h, _handles(oracles), len(oracles), C.byref(instance.struct), _ptr(a),
                                       _ptr(out)))
    return out
