"""Host mirror of the prover steps around the commitments (plonky2 plonk/prover.rs): the circuit description that
`compute_quotient_polys` reads from `CommonCircuitData`, and the device quotient evaluation behind
bsx_quotient_values / bsx_compute_quotient_polys.  Also a synthetic circuit of the built-in gates with a
satisfying witness (the plonky2x circuits of BlobstreamX are not in the reference mount), used by tests and bench."""
import ctypes as C
from dataclasses import dataclass, field
from typing import List

import numpy as np

from .binding import BsxError, PolynomialBatch, Timing, lib, _np, _ptr

P = 0xFFFFFFFF00000001
GATE_NOOP, GATE_CONSTANT, GATE_PUBLIC_INPUT, GATE_ARITHMETIC = 0, 1, 2, 3
# gates/poseidon.rs PoseidonGate (135 wires, 123 constraints), gates/arithmetic_extension.rs ArithmeticExtensionGate<2>,
# gates/multiplication_extension.rs MulExtensionGate<2> (param = num_ops), gates/base_sum.rs BaseSumGate<2> (param = num_limbs)
GATE_POSEIDON, GATE_ARITHMETIC_EXTENSION, GATE_MUL_EXTENSION, GATE_BASE_SUM = 4, 5, 6, 7
# gates/exponentiation.rs ExponentiationGate (param = num_power_bits), gates/reducing.rs ReducingGate<2> and
# gates/reducing_extension.rs ReducingExtensionGate<2> (param = num_coeffs), gates/random_access.rs RandomAccessGate<2>
# (param = bits | num_copies << 8 | num_extra_constants << 16), gates/poseidon_mds.rs PoseidonMdsGate<2>
GATE_EXPONENTIATION, GATE_REDUCING, GATE_REDUCING_EXTENSION, GATE_RANDOM_ACCESS, GATE_POSEIDON_MDS = 8, 9, 10, 11, 12
GATE_COSET_INTERPOLATION = 13  # gates/coset_interpolation.rs CosetInterpolationGate<2>: param = subgroup_bits | degree << 8
POSEIDON_GATE_WIRES, POSEIDON_GATE_CONSTRAINTS = 135, 123


class _Gate(C.Structure):
    _fields_ = [("kind", C.c_int32), ("param", C.c_int32), ("row", C.c_int32), ("selector_index", C.c_int32),
                ("group_start", C.c_int32), ("group_end", C.c_int32)]


class _Circuit(C.Structure):
    _fields_ = [("degree_bits", C.c_int32), ("rate_bits", C.c_int32), ("quotient_degree_factor", C.c_int32),
                ("num_challenges", C.c_int32), ("num_wires", C.c_int32), ("num_routed_wires", C.c_int32),
                ("num_constants", C.c_int32), ("num_selectors", C.c_int32), ("num_partial_products", C.c_int32),
                ("num_gate_constraints", C.c_int32), ("n_gates", C.c_int32), ("gates", C.c_void_p), ("k_is", C.c_void_p)]


@dataclass
class Gate:
    """One entry of `common_data.gates` with its share of `selectors_info`."""
    kind: int
    param: int = 0
    row: int = 0
    selector_index: int = 0
    group_start: int = 0
    group_end: int = 1

    def num_constraints(self):
        return {GATE_NOOP: 0, GATE_CONSTANT: self.param, GATE_PUBLIC_INPUT: 4, GATE_ARITHMETIC: self.param,
                GATE_POSEIDON: POSEIDON_GATE_CONSTRAINTS, GATE_ARITHMETIC_EXTENSION: 2 * self.param,
                GATE_MUL_EXTENSION: 2 * self.param, GATE_BASE_SUM: 1 + self.param,
                GATE_EXPONENTIATION: self.param + 1, GATE_REDUCING: 2 * self.param, GATE_REDUCING_EXTENSION: 2 * self.param,
                GATE_RANDOM_ACCESS: ((self.param >> 8) & 0xFF) * ((self.param & 0xFF) + 2) + (self.param >> 16),
                GATE_POSEIDON_MDS: 24,
                GATE_COSET_INTERPOLATION: 2 * (2 + 2 * (((1 << (self.param & 0xFF)) - 2) // max(1, (self.param >> 8) - 1)))}[self.kind]


@dataclass
class CircuitDescription:
    """The fields of `CommonCircuitData` / `CircuitConfig` the quotient evaluation reads.  Defaults are
    `CircuitConfig::standard_recursion_config()`."""
    degree_bits: int
    gates: List[Gate]
    rate_bits: int = 3
    quotient_degree_factor: int = 8
    num_challenges: int = 2
    num_wires: int = 135
    num_routed_wires: int = 80
    num_constants: int = 3          # selectors first, then the gates' constants
    num_selectors: int = 1
    k_is: np.ndarray = field(default=None)

    def __post_init__(self):
        if self.k_is is None:  # plonk/plonk_common.rs get_unique_coset_shifts: g^j, g = 7
            self.k_is = np.array([pow(7, j, P) for j in range(self.num_routed_wires)], dtype=np.uint64)

    @property
    def num_partial_products(self):
        return -(-self.num_routed_wires // self.quotient_degree_factor) - 1

    @property
    def num_gate_constraints(self):
        return max([g.num_constraints() for g in self.gates] + [0])

    @property
    def quotient_degree_bits(self):
        return max(self.quotient_degree_factor - 1, 0).bit_length()

    def as_dict(self):
        """The oracle's view of the same description (oracle.quotient_values)."""
        return {"degree_bits": self.degree_bits, "rate_bits": self.rate_bits,
                "quotient_degree_factor": self.quotient_degree_factor, "num_challenges": self.num_challenges,
                "num_wires": self.num_wires, "num_routed_wires": self.num_routed_wires,
                "num_constants": self.num_constants, "num_selectors": self.num_selectors,
                "num_partial_products": self.num_partial_products, "num_gate_constraints": self.num_gate_constraints,
                "gates": [(g.kind, g.param, g.row, g.selector_index, g.group_start, g.group_end) for g in self.gates],
                "k_is": self.k_is}

    def _struct(self):
        gates = (_Gate * max(len(self.gates), 1))(*[_Gate(g.kind, g.param, g.row, g.selector_index, g.group_start,
                                                          g.group_end) for g in self.gates])
        k_is = _np(self.k_is)
        c = _Circuit(self.degree_bits, self.rate_bits, self.quotient_degree_factor, self.num_challenges, self.num_wires,
                     self.num_routed_wires, self.num_constants, self.num_selectors, self.num_partial_products,
                     self.num_gate_constraints, len(self.gates), C.cast(gates, C.c_void_p), k_is.ctypes.data)
        return c, (gates, k_is)


def _declare():
    L = lib()
    if getattr(L, "_bsx_plonk_declared", False):
        return L
    vp, i32 = C.c_void_p, C.c_int32
    L.bsx_quotient_values.restype = i32
    L.bsx_quotient_values.argtypes = [vp, C.POINTER(_Circuit), vp, vp, vp, vp, vp, vp, vp, vp, i32]
    L.bsx_compute_quotient_polys.restype = i32
    L.bsx_compute_quotient_polys.argtypes = [vp, C.POINTER(_Circuit), vp, vp, vp, vp, vp, vp, vp, i32, i32,
                                             C.POINTER(vp), C.POINTER(Timing), C.POINTER(C.c_float)]
    L._bsx_plonk_declared = True
    return L


def _challenges(circuit, public_inputs_hash, betas, gammas, alphas):
    pih = _np(public_inputs_hash if public_inputs_hash is not None else np.zeros(4, dtype=np.uint64)).reshape(4)
    b, g, a = _np(betas).reshape(-1), _np(gammas).reshape(-1), _np(alphas).reshape(-1)
    if not (b.size == g.size == a.size == circuit.num_challenges):
        raise BsxError("betas / gammas / alphas must each hold num_challenges elements")
    return pih, b, g, a


def quotient_values(ctx, circuit, constants_sigmas, wires, zs_partial_products, public_inputs_hash, betas, gammas, alphas,
                    out_device_ptr=None):
    """`quotient_values` of plonk/prover.rs compute_quotient_polys, transposed: (num_challenges, 2^(degree_bits + q))
    values on the coset, before `coset_ifft`.  The commitments are device-resident PolynomialBatch objects."""
    c, keep = circuit._struct()
    pih, b, g, a = _challenges(circuit, public_inputs_hash, betas, gammas, alphas)
    out = None
    if out_device_ptr is None:
        out = np.zeros((circuit.num_challenges, 1 << (circuit.degree_bits + circuit.quotient_degree_bits)), dtype=np.uint64)
    ctx.check(_declare().bsx_quotient_values(
        ctx._h, C.byref(c), constants_sigmas._h, wires._h, zs_partial_products._h, _ptr(pih), _ptr(b), _ptr(g), _ptr(a),
        _ptr(out) if out is not None else C.c_void_p(int(out_device_ptr)), 0 if out is not None else 1))
    return out


def compute_quotient_polys(ctx, circuit, constants_sigmas, wires, zs_partial_products, public_inputs_hash, betas, gammas,
                           alphas, cap_height, blinding=False, timing=False):
    """plonk/prover.rs: `compute_quotient_polys(..)` followed by prove()'s `trim_to_len`, `.chunks(degree)` and
    `PolynomialBatch::from_coeffs(all_quotient_poly_chunks, ..)` -> the quotient commitment.  Raises where upstream
    panics (a witness that does not satisfy the circuit: "Quotient has failed").  The batch gets `.eval_ms`."""
    c, keep = circuit._struct()
    pih, b, g, a = _challenges(circuit, public_inputs_hash, betas, gammas, alphas)
    h = C.c_void_p()
    t = Timing() if timing else None
    ms = C.c_float()
    ctx.check(_declare().bsx_compute_quotient_polys(
        ctx._h, C.byref(c), constants_sigmas._h, wires._h, zs_partial_products._h, _ptr(pih), _ptr(b), _ptr(g), _ptr(a),
        int(cap_height), 1 if blinding else 0, C.byref(h), C.byref(t) if t is not None else None, C.byref(ms)))
    batch = PolynomialBatch(ctx, h, t)
    batch.eval_ms = float(ms.value)
    return batch


# ---------------------------------------------------------------------------------------------
def _mulmod(a, b):
    """elementwise a * b mod p on uint64 arrays (python ints under the hood: host-side test data only)."""
    return np.array([(int(x) * int(y)) % P for x, y in zip(np.ravel(a), np.ravel(b))], dtype=np.uint64).reshape(np.shape(a))


@dataclass
class SyntheticCircuit:
    circuit: CircuitDescription
    constants: np.ndarray       # (num_constants, n): selector column, then const_0, const_1
    sigmas: np.ndarray          # (num_routed_wires, n): sigma_j(omega^r) = k_j' omega^r'
    wires: np.ndarray           # (num_wires, n): a witness that satisfies every gate and every copy constraint
    public_inputs_hash: np.ndarray


UNUSED_SELECTOR = 0xFFFFFFFF  # gates/selectors.rs: what a selector column holds on rows of another group


def _ext_mul(a, b):
    """(a0 + a1 X)(b0 + b1 X) in F_p[X] / (X^2 - 7) on python ints"""
    return (a[0] * b[0] + 7 * a[1] * b[1]) % P, (a[0] * b[1] + a[1] * b[0]) % P


def _all_gates(num_wires, num_routed_wires, num_ops, with_poseidon):
    """The gate table of gate_set "all": every kind that fits the wire counts, in selector groups that keep
    gate degree + group size <= 9 (gates/selectors.rs for quotient_degree_factor 8)."""
    ra_bits = 4
    while ra_bits > 1 and (2 + (1 << ra_bits)) > min(num_routed_wires, num_wires - ra_bits):
        ra_bits -= 1
    vec = 1 << ra_bits
    ra_copies = min(num_routed_wires // (2 + vec), num_wires // (2 + vec + ra_bits))
    ra_extra = max(0, min(num_routed_wires - (2 + vec) * ra_copies, 2, num_wires - (2 + vec + ra_bits) * ra_copies))
    # CosetInterpolationGate::with_max_degree(subgroup_bits, 8): the smallest degree with as few intermediates as degree 8
    ci_bits = 4 if (num_wires >= 47 and num_routed_wires >= 37) else 2 if (num_wires >= 15 and num_routed_wires >= 13) else 0
    n_int = lambda d: ((1 << ci_bits) - 2) // (d - 1)
    ci_degree = min((d for d in range(2, min(8, 1 << ci_bits) + 1) if n_int(d) == n_int(min(8, 1 << ci_bits))), default=0) if ci_bits else 0
    groups = [
        [(GATE_NOOP, 0), (GATE_CONSTANT, 2), (GATE_PUBLIC_INPUT, 0), (GATE_POSEIDON_MDS, 0),
         (GATE_BASE_SUM, min(32, num_wires - 1)), (GATE_REDUCING, min(num_routed_wires - 6, (num_wires - 4) // 3))],
        [(GATE_REDUCING_EXTENSION, min((num_routed_wires - 6) // 2, (num_wires - 4) // 4)), (GATE_ARITHMETIC, num_ops),
         (GATE_ARITHMETIC_EXTENSION, num_routed_wires // 8), (GATE_MUL_EXTENSION, num_routed_wires // 6),
         (GATE_EXPONENTIATION, min(num_routed_wires - 3, (num_wires - 2) // 2))],
        [(GATE_RANDOM_ACCESS, ra_bits | ra_copies << 8 | ra_extra << 16), (GATE_COSET_INTERPOLATION, ci_bits | ci_degree << 8)],
        [(GATE_POSEIDON, 0)]]
    fits = lambda k, p: not ((k == GATE_POSEIDON and not (with_poseidon and num_wires >= POSEIDON_GATE_WIRES)) or
                             (k == GATE_POSEIDON_MDS and num_wires < 48) or (k == GATE_RANDOM_ACCESS and ra_copies < 1) or
                             (k in (GATE_REDUCING, GATE_REDUCING_EXTENSION, GATE_EXPONENTIATION) and p < 1) or
                             (k == GATE_COSET_INTERPOLATION and ci_bits == 0))
    gates, row = [], 0
    for sel, grp in enumerate(groups):
        grp = [(k, p) for k, p in grp if fits(k, p)]
        for k, p in grp:
            gates.append(Gate(k, p, row + len([1 for g in gates if g.selector_index == sel]), sel, row, row + len(grp)))
        row += len(grp)
    return gates


def synthetic_circuit(degree_bits, seed=1, n_copies=None, num_wires=135, num_routed_wires=80, num_ops=20,
                      quotient_degree_factor=8, rate_bits=3, num_challenges=2, public_inputs_hash=None, gate_set="basic",
                      poseidon_fill=None, max_poseidon_rows=1 << 12):
    """A circuit with random constants, `n_copies` copy constraints between free arithmetic inputs, and a witness that
    satisfies all of it.  gate_set "basic": NoopGate / ConstantGate / PublicInputGate / ArithmeticGate rows in one selector
    group.  "all" (degree_bits >= 6, quotient_degree_factor 8): four selector groups as gates/selectors.rs would form them
    for max degree 9 -- [Noop, Constant, PublicInput, PoseidonMds, BaseSum<2>, Reducing], [ReducingExtension, Arithmetic,
    ArithmeticExtension, MulExtension, Exponentiation], [RandomAccess, CosetInterpolation] and [Poseidon] -- with rows of
    every kind;
    `poseidon_fill(rows)` completes (k, 135) PoseidonGate rows whose inputs and swap wire are set
    (Context.poseidon_gate_witness on the device; the Poseidon gate is left out when it is None or the circuit has fewer
    than 135 wires)."""
    n = 1 << degree_bits
    rng = np.random.default_rng(seed)
    rand = lambda shape: rng.integers(0, P, size=shape, dtype=np.uint64)
    if gate_set == "all":
        if degree_bits < 6 or quotient_degree_factor != 8:
            raise ValueError('gate_set "all" wants degree_bits >= 6 and quotient_degree_factor 8')
        gates = _all_gates(num_wires, num_routed_wires, num_ops, poseidon_fill is not None)
        n_sel = 4
        circ = CircuitDescription(degree_bits, gates, rate_bits=rate_bits, quotient_degree_factor=quotient_degree_factor,
                                  num_challenges=num_challenges, num_wires=num_wires, num_routed_wires=num_routed_wires,
                                  num_constants=n_sel + 2, num_selectors=n_sel)
        kind_row = np.full(n, GATE_ARITHMETIC, dtype=np.int64)   # arithmetic rows ...
        kind_row[0] = GATE_PUBLIC_INPUT                           # ... one public-input row,
        at = 1
        for g in gates:                                           # a run of rows of every other kind,
            if g.kind in (GATE_NOOP, GATE_PUBLIC_INPUT, GATE_ARITHMETIC):
                continue
            k = max(1, n // 32)
            if g.kind == GATE_POSEIDON:
                k = min(max(1, n // 8), max_poseidon_rows)
            kind_row[at:at + k] = g.kind
            at += k
        kind_row[n - max(1, n // 32):] = GATE_NOOP                # and padding rows at the end
    else:
        gates = [Gate(GATE_NOOP, 0, 0, 0, 0, 4), Gate(GATE_CONSTANT, 2, 1, 0, 0, 4), Gate(GATE_PUBLIC_INPUT, 0, 2, 0, 0, 4),
                 Gate(GATE_ARITHMETIC, num_ops, 3, 0, 0, 4)]
        n_sel = 1
        circ = CircuitDescription(degree_bits, gates, rate_bits=rate_bits, quotient_degree_factor=quotient_degree_factor,
                                  num_challenges=num_challenges, num_wires=num_wires, num_routed_wires=num_routed_wires)
        kind_row = np.full(n, GATE_ARITHMETIC, dtype=np.int64)    # arithmetic rows ...
        kind_row[0] = GATE_PUBLIC_INPUT                           # ... one public-input row,
        n_const = max(1, n // 16) if n > 2 else 0
        kind_row[1:1 + n_const] = GATE_CONSTANT                   # some constant rows,
        if n > 4:
            kind_row[n - max(1, n // 32):] = GATE_NOOP            # and padding rows at the end
    by_kind = {g.kind: g for g in gates}
    gate_idx = np.array([by_kind[int(k)].row for k in kind_row])
    gate_sel = np.array([by_kind[int(k)].selector_index for k in kind_row])
    sel_cols = [np.where(gate_sel == t, gate_idx, UNUSED_SELECTOR).astype(np.uint64) for t in range(n_sel)]
    constants = np.stack(sel_cols + [rand(n), rand(n)])
    c0, c1 = constants[n_sel], constants[n_sel + 1]
    rows_of = lambda kind: np.nonzero(kind_row == kind)[0]
    param_of = lambda kind: by_kind[kind].param
    wires = rand((num_wires, n))
    pih = rand(4) if public_inputs_hash is None else np.array(public_inputs_hash, dtype=np.uint64).reshape(4).copy()
    # copy constraints: disjoint pairs of free cells (multiplicands / addends of arithmetic rows)
    arith_rows = rows_of(GATE_ARITHMETIC)
    free = [(4 * i + t, r) for r in arith_rows[: 4096] for i in range(num_ops) for t in range(3)]
    if n_copies is None:
        n_copies = min(len(free) // 2, 4 * n)
    pick = rng.permutation(len(free))[: 2 * n_copies]
    ident = np.arange(num_routed_wires * n, dtype=np.int64).reshape(num_routed_wires, n)
    perm = ident.copy()
    for a, b in zip(pick[0::2], pick[1::2]):
        (ja, ra), (jb, rb) = free[a], free[b]
        wires[jb, rb] = wires[ja, ra]
        perm[ja, ra], perm[jb, rb] = ident[jb, rb], ident[ja, ra]
    # satisfy the gates
    w = pow(1753635133440165772, 1 << (32 - degree_bits), P)  # plonky2 primitive_root_of_unity(degree_bits)
    c_rows = rows_of(GATE_CONSTANT)
    wires[0, c_rows], wires[1, c_rows] = c0[c_rows], c1[c_rows]
    wires[0:4, 0] = pih
    for i in range(num_ops):
        m = _mulmod(_mulmod(wires[4 * i, arith_rows], wires[4 * i + 1, arith_rows]), c0[arith_rows])
        a = _mulmod(wires[4 * i + 2, arith_rows], c1[arith_rows])
        wires[4 * i + 3, arith_rows] = np.array([(int(x) + int(y)) % P for x, y in zip(m, a)], dtype=np.uint64)
    if gate_set == "all":
        _fill_gate_rows(wires, rows_of, param_of, by_kind, c0, c1, rng, poseidon_fill)
    # sigma polynomials' values: cell (j', r') -> k_j' * omega^r'
    omegas = np.empty(n, dtype=object)
    acc = 1
    for r in range(n):
        omegas[r] = acc
        acc = acc * w % P
    k = [int(x) for x in circ.k_is]
    sig = np.empty((num_routed_wires, n), dtype=np.uint64)
    flat = perm.reshape(-1)
    jj, rr = flat // n, flat % n
    sig.reshape(-1)[:] = np.array([k[j] * omegas[r] % P for j, r in zip(jj, rr)], dtype=np.uint64)
    return SyntheticCircuit(circ, constants, sig, wires, pih)


MDS_CIRC = [17, 15, 41, 16, 2, 28, 13, 13, 39, 18, 34, 20]  # hash/poseidon_goldilocks.rs MDS_MATRIX_CIRC; DIAG = [8, 0, ...]


def _fill_gate_rows(wires, rows_of, param_of, by_kind, c0, c1, rng, poseidon_fill):
    """What the gates' generators (SimpleGenerator::run_once of each gate file) compute, on python ints."""
    W = lambda j, r: int(wires[j, r])
    if GATE_BASE_SUM in by_kind:                     # wire 0 = sum of 2^i * limb_i, limbs are bits
        n_limbs = param_of(GATE_BASE_SUM)
        for r in rows_of(GATE_BASE_SUM):
            bits = rng.integers(0, 2, size=n_limbs, dtype=np.uint64)
            wires[1:1 + n_limbs, r] = bits
            wires[0, r] = sum(int(b) << i for i, b in enumerate(bits)) % P
    for kind, per in ((GATE_ARITHMETIC_EXTENSION, 8), (GATE_MUL_EXTENSION, 6)):  # output = m0 m1 const_0 (+ addend const_1) in F_p^2
        for r in rows_of(kind) if kind in by_kind else ():
            k0, k1 = int(c0[r]), int(c1[r])
            for i in range(param_of(kind)):
                q = [W(per * i + t, r) for t in range(per)]
                pr = _ext_mul(q[0:2], q[2:4])
                out = [pr[d] * k0 % P for d in range(2)]
                if kind == GATE_ARITHMETIC_EXTENSION:
                    out = [(out[d] + q[4 + d] * k1) % P for d in range(2)]
                wires[per * i + per - 2:per * i + per, r] = out
    if GATE_EXPONENTIATION in by_kind:               # base 0, power bits 1.. (little-endian), output 1 + n, intermediates 2 + n..
        nb = param_of(GATE_EXPONENTIATION)
        for r in rows_of(GATE_EXPONENTIATION):
            bits = rng.integers(0, 2, size=nb, dtype=np.uint64)
            wires[1:1 + nb, r] = bits
            base, acc = W(0, r), 1
            for i in range(nb):
                acc = acc * acc % P * (base if bits[nb - 1 - i] else 1) % P
                wires[2 + nb + i, r] = acc
            wires[1 + nb, r] = acc
    for kind, cw in ((GATE_REDUCING, 1), (GATE_REDUCING_EXTENSION, 2)):  # acc_i = acc_{i-1} alpha + coeff_i, last one = output
        for r in rows_of(kind) if kind in by_kind else ():
            nco = param_of(kind)
            alpha, acc = (W(2, r), W(3, r)), (W(4, r), W(5, r))
            for i in range(nco):
                t = _ext_mul(acc, alpha)
                acc = ((t[0] + W(6 + cw * i, r)) % P, (t[1] + (W(6 + 2 * i + 1, r) if cw == 2 else 0)) % P)
                at = 0 if i == nco - 1 else 6 + nco * cw + 2 * i
                wires[at:at + 2, r] = acc
    if GATE_RANDOM_ACCESS in by_kind:                # per copy: access_index, claimed_element, 2^bits items; bits after the routed wires
        p = param_of(GATE_RANDOM_ACCESS)
        bits, copies, extra = p & 0xFF, (p >> 8) & 0xFF, p >> 16
        vec = 1 << bits
        routed = (2 + vec) * copies + extra
        for r in rows_of(GATE_RANDOM_ACCESS):
            for c in range(copies):
                idx = int(rng.integers(0, vec))
                q = (2 + vec) * c
                wires[q, r] = idx
                wires[q + 1, r] = wires[q + 2 + idx, r]
                wires[routed + c * bits:routed + (c + 1) * bits, r] = [(idx >> i) & 1 for i in range(bits)]
            for i in range(extra):
                wires[(2 + vec) * copies + i, r] = (c0, c1)[i][r]
    if GATE_COSET_INTERPOLATION in by_kind:          # shift 0, values 1 + 2 i, evaluation point, value, intermediates, shifted point
        p = param_of(GATE_COSET_INTERPOLATION)
        bits, degree = p & 0xFF, p >> 8
        npts = 1 << bits
        ni = (npts - 2) // (degree - 1)
        ep, ev, im = 1 + 2 * npts, 3 + 2 * npts, 5 + 2 * npts
        sp = im + 4 * ni
        gen = pow(1753635133440165772, 1 << (32 - bits), P)
        dom = [pow(gen, i, P) for i in range(npts)]
        wts = []
        for i in range(npts):
            pr = 1
            for j in range(npts):
                if j != i:
                    pr = pr * (dom[i] - dom[j]) % P
            wts.append(pow(pr, P - 2, P))
        for r in rows_of(GATE_COSET_INTERPOLATION):
            shift = W(0, r) or 1
            wires[0, r] = shift
            x = (W(sp, r), W(sp + 1, r))
            wires[ep, r], wires[ep + 1, r] = x[0] * shift % P, x[1] * shift % P
            ev_, pr_ = (0, 0), (1, 0)
            first, last = 0, degree
            for i in range(ni + 1):
                for q in range(first, last):   # partial_interpolate
                    val = (W(1 + 2 * q, r) * wts[q] % P, W(2 + 2 * q, r) * wts[q] % P)
                    term = ((x[0] - dom[q]) % P, x[1])
                    a, b = _ext_mul(ev_, term), _ext_mul(val, pr_)
                    ev_, pr_ = ((a[0] + b[0]) % P, (a[1] + b[1]) % P), _ext_mul(pr_, term)
                if i == ni:
                    break
                wires[im + 2 * i:im + 2 * i + 2, r] = ev_
                wires[im + 2 * (ni + i):im + 2 * (ni + i) + 2, r] = pr_
                first = 1 + (degree - 1) * (i + 1)
                last = min(first + degree - 1, npts)
            wires[ev:ev + 2, r] = ev_
    if GATE_POSEIDON_MDS in by_kind:                 # outputs (wires 24..47) = MDS * inputs (wires 0..23), component by component
        for r in rows_of(GATE_POSEIDON_MDS):
            for d in range(2):
                x = [W(2 * i + d, r) for i in range(12)]
                for o in range(12):
                    wires[24 + 2 * o + d, r] = (sum(MDS_CIRC[i] * x[(i + o) % 12] for i in range(12)) + (8 * x[0] if o == 0 else 0)) % P
    p_rows = rows_of(GATE_POSEIDON) if GATE_POSEIDON in by_kind else np.zeros(0, dtype=np.int64)
    if p_rows.size:
        rows = np.ascontiguousarray(wires[:POSEIDON_GATE_WIRES, p_rows].T)
        rows[:, 24] = rng.integers(0, 2, size=p_rows.size, dtype=np.uint64)  # swap
        rows = _np(poseidon_fill(rows)).reshape(p_rows.size, POSEIDON_GATE_WIRES)
        wires[:POSEIDON_GATE_WIRES, p_rows] = rows.T


# ---------------------------------------------------------------------------------------------
def hash_no_pad(ctx, elems):
    """plonky2 hash/hashing.rs `hash_n_to_hash_no_pad` (overwrite-mode sponge, rate 8) with the permutations on the
    device: host-side glue for the few small hashes of prove() (public inputs, circuit digest)."""
    state = np.zeros(12, dtype=np.uint64)
    e = _np(elems).reshape(-1)
    for off in range(0, e.size, 8):
        chunk = e[off:off + 8]
        state[:chunk.size] = chunk
        state = ctx.poseidon_permute(state)
    return state[:4].copy()


def circuit_digest(ctx, constants_sigmas_cap, degree_bits):
    """plonk/circuit_builder.rs: `circuit_digest = hash_no_pad(constants_sigmas_cap.flatten() ++ [degree_bits])` (the
    revisions BlobstreamX could pin differ in what else they append; the synthetic prove uses this form)."""
    parts = np.concatenate([_np(constants_sigmas_cap).reshape(-1), np.array([degree_bits], dtype=np.uint64)])
    return hash_no_pad(ctx, parts)


def prove(ctx, sc, fri_config, public_inputs, constants_sigmas=None):
    """plonk/prover.rs `prove()` for a SyntheticCircuit, commitment phase on the device end to end: wires commitment,
    challenges, Zs and partial products (device), their commitment, the quotient evaluated from the resident LDEs and
    committed, the opening set at zeta and g zeta, prove_openings.  The witness is given (sc.wires): witness generation is
    upstream's sequential generator pass and out of scope.  Returns (ProofWithPublicInputs, dict of what a verifier needs
    besides the proof: circuit digest, constants_sigmas cap, FRI parameters, oracle widths)."""
    from .fri import Challenger, FriInstanceInfo, prove_openings
    from .proof import OpeningSet, Proof, ProofWithPublicInputs
    c = sc.circuit
    n, nc = 1 << c.degree_bits, c.num_challenges
    fp = fri_config.fri_params(c.degree_bits)
    own_cs = constants_sigmas is None
    if own_cs:  # circuit data: committed at build time upstream
        constants_sigmas = PolynomialBatch.from_values(ctx, np.concatenate([sc.constants, sc.sigmas]), fri_config.rate_bits,
                                                       False, fri_config.cap_height)
    digest = circuit_digest(ctx, constants_sigmas.cap, c.degree_bits)
    pih = hash_no_pad(ctx, public_inputs)
    if not (pih == _np(sc.public_inputs_hash)).all():
        raise BsxError("the witness was built for another public-input hash")
    ch = Challenger(ctx)
    ch.observe_hash(digest)
    ch.observe_hash(pih)
    wires = PolynomialBatch.from_values(ctx, sc.wires, fri_config.rate_bits, False, fri_config.cap_height)
    ch.observe_cap(wires.cap)
    betas, gammas = ch.get_n_challenges(nc), ch.get_n_challenges(nc)
    zs_vals = ctx.partial_products_and_zs(sc.wires[:c.num_routed_wires], sc.sigmas, c.k_is, betas, gammas,
                                          c.quotient_degree_factor)
    zs = PolynomialBatch.from_values(ctx, zs_vals, fri_config.rate_bits, False, fri_config.cap_height)
    ch.observe_cap(zs.cap)
    alphas = ch.get_n_challenges(nc)
    quotient = compute_quotient_polys(ctx, c, constants_sigmas, wires, zs, pih, betas, gammas, alphas, fri_config.cap_height)
    ch.observe_cap(quotient.cap)
    zeta = ch.get_extension_challenge()
    g = pow(1753635133440165772, 1 << (32 - c.degree_bits), P)  # F::primitive_root_of_unity(degree_bits)
    openings = OpeningSet.new(zeta, g, constants_sigmas, wires, zs, quotient, c.num_constants, nc)
    for batch in openings.to_fri_openings():  # challenger.observe_openings
        ch.observe_extension_elements(batch.reshape(-1))
    widths = [c.num_constants + c.num_routed_wires, c.num_wires, nc * (1 + c.num_partial_products),
              nc * c.quotient_degree_factor]
    gz = np.array([int(zeta[0]) * g % P, int(zeta[1]) * g % P], dtype=np.uint64)
    instance = FriInstanceInfo(np.stack([zeta, gz]), [[(o, j) for o, w in enumerate(widths) for j in range(w)],
                                                      [(2, j) for j in range(nc)]])
    oracles = [constants_sigmas, wires, zs, quotient]
    opening_proof = prove_openings(instance, oracles, ch, fp)
    pw = ProofWithPublicInputs(Proof(wires.cap, zs.cap, quotient.cap, openings, opening_proof), _np(public_inputs).reshape(-1))
    aux = {"circuit_digest": digest, "constants_sigmas_cap": constants_sigmas.cap, "fri_params": fp, "widths": widths}
    for b in oracles[1:] + ([constants_sigmas] if own_cs else []):
        b.close()
    return pw, aux
