"""ctypes binding of the CPU oracle (oracle/liboracle.so).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and the
cpu_baseline / --impl reference legs of bench.py.  The product package never imports it.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_DIR = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_DIR, "liboracle.so")
P = 0xFFFFFFFF00000001
u64p = np.ctypeslib.ndpointer(dtype=np.uint64, flags="C_CONTIGUOUS")


def build(force=False):
    srcs = [os.path.join(_DIR, f) for f in ("plonky2_oracle.c", "plonky2_fri_oracle.c", "plonky2_plonk_oracle.c", "plonky2_oracle.h",
                                              "poseidon_rc.h", "Makefile")]
    if force or not os.path.exists(_SO) or any(os.path.getmtime(s) > os.path.getmtime(_SO) for s in srcs):
        subprocess.check_call(["make", "-C", _DIR, "-s"])
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_SO)
        L = _lib
        for name in ("orc_add", "orc_sub", "orc_mul", "orc_pow"):
            getattr(L, name).restype = C.c_uint64
            getattr(L, name).argtypes = [C.c_uint64, C.c_uint64]
        L.orc_inv.restype = C.c_uint64
        L.orc_inv.argtypes = [C.c_uint64]
        L.orc_root_of_unity.restype = C.c_uint64
        L.orc_root_of_unity.argtypes = [C.c_int]
        L.orc_reverse_bits.restype = C.c_uint64
        L.orc_reverse_bits.argtypes = [C.c_uint64, C.c_int]
        L.orc_round_constants.restype = C.POINTER(C.c_uint64)
        L.orc_num_threads.restype = C.c_int
    return _lib


def _a(x):
    return np.ascontiguousarray(x, dtype=np.uint64)


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def fft(a):
    a = _a(a).copy()
    lib().orc_fft(_ptr(a), C.c_int(int(np.log2(a.size))))
    return a


def ifft(a):
    a = _a(a).copy()
    lib().orc_ifft(_ptr(a), C.c_int(int(np.log2(a.size))))
    return a


def coset_fft(a, shift=7):
    a = _a(a).copy()
    lib().orc_coset_fft(_ptr(a), C.c_int(int(np.log2(a.size))), C.c_uint64(shift))
    return a


def coset_ifft(a, shift=7):
    """PolynomialValues::coset_ifft (field/src/polynomial/mod.rs)"""
    a = _a(a).copy()
    lib().orc_coset_ifft(_ptr(a), C.c_int(int(np.log2(a.size))), C.c_uint64(shift))
    return a


def quotient_chunks(quotient_values, degree_log, quotient_degree_factor):
    """plonk/prover.rs: `values.coset_ifft(F::coset_shift())` per quotient polynomial, `trim_to_len(quotient_degree)`
    (raises where upstream's expect() panics), `.chunks(degree)`: (n_polys * factor, 2^degree_log) coefficient rows."""
    n = 1 << degree_log
    out = []
    for v in _a(quotient_values):
        c = coset_ifft(v)
        keep = quotient_degree_factor * n
        if c[keep:].any():
            raise ValueError("Quotient has failed, the vanishing polynomial is not divisible by Z_H")
        out.extend(c[j * n:(j + 1) * n] for j in range(quotient_degree_factor))
    return np.stack(out)


def partial_products_and_zs(wires, sigmas, k_is, betas, gammas, max_degree):
    """plonk/prover.rs all_wires_permutation_partial_products followed by prove()'s reordering ("Z is expected at the
    front of our batch"): rows Z_0 .. Z_{c-1}, then every challenge's partial products."""
    wires, sigmas, k_is = _a(wires), _a(sigmas), _a(k_is)
    n_routed, n = wires.shape
    n_chunks = -(-n_routed // max_degree)
    zs, pps = [], []
    for beta, gamma in zip(betas, gammas):
        z = np.empty(n, dtype=np.uint64)
        pp = np.empty((n_chunks - 1, n), dtype=np.uint64)
        rc = lib().orc_partial_products_and_z(_ptr(wires), _ptr(sigmas), C.c_int(int(np.log2(n))), C.c_int(n_routed),
                                              C.c_int(max_degree), _ptr(k_is), C.c_uint64(int(beta) % P),
                                              C.c_uint64(int(gamma) % P), _ptr(z), _ptr(pp))
        if rc != 0:
            raise ZeroDivisionError("Tried to invert zero")
        zs.append(z)
        pps.append(pp)
    return np.concatenate([np.stack(zs)] + pps, axis=0)


def lde(coeffs, rate_bits, shift=7):
    coeffs = _a(coeffs)
    out = np.empty(coeffs.size << rate_bits, dtype=np.uint64)
    lib().orc_lde(_ptr(coeffs), C.c_int(int(np.log2(coeffs.size))), C.c_int(rate_bits), C.c_uint64(shift), _ptr(out))
    return out


def poseidon(states):
    """Permute one (12,) state or a batch (n, 12)."""
    s = _a(states).copy()
    flat = s.reshape(-1, 12)
    for i in range(flat.shape[0]):
        lib().orc_poseidon(C.c_void_p(flat[i].ctypes.data))
    return s


def hash_no_pad(x):
    x = _a(x)
    out = np.empty(4, dtype=np.uint64)
    lib().orc_hash_no_pad(_ptr(x), C.c_size_t(x.size), _ptr(out))
    return out


def hash_or_noop(x):
    x = _a(x)
    out = np.empty(4, dtype=np.uint64)
    lib().orc_hash_or_noop(_ptr(x), C.c_size_t(x.size), _ptr(out))
    return out


def two_to_one(l, r):
    l, r = _a(l), _a(r)
    out = np.empty(4, dtype=np.uint64)
    lib().orc_two_to_one(_ptr(l), _ptr(r), _ptr(out))
    return out


def n_digest_words(log_leaves, cap_height):
    return 4 * 2 * ((1 << log_leaves) - (1 << cap_height))


def merkle_build(leaves, cap_height):
    """leaves: (n_leaves, leaf_len) row-major.  Returns (digests level-order (k,4), cap (2^cap,4))."""
    leaves = _a(leaves)
    n, w = leaves.shape
    log_leaves = int(np.log2(n))
    assert 1 << log_leaves == n, "leaf count must be a power of two"
    digests = np.zeros(n_digest_words(log_leaves, cap_height) if cap_height <= log_leaves else 0, dtype=np.uint64)
    cap = np.zeros(4 << min(cap_height, 62), dtype=np.uint64) if cap_height <= log_leaves else np.zeros(4, np.uint64)
    rc = lib().orc_merkle_build(_ptr(leaves), C.c_int(log_leaves), C.c_size_t(w), C.c_int(cap_height),
                                _ptr(digests) if digests.size else None, _ptr(cap))
    if rc != 0:
        raise ValueError("cap_height %d > log2(leaves) %d" % (cap_height, log_leaves))
    return digests.reshape(-1, 4), cap.reshape(-1, 4)


def merkle_prove(digests, log_leaves, cap_height, leaf_index):
    digests = _a(digests)
    sib = np.zeros(4 * (log_leaves - cap_height), dtype=np.uint64)
    rc = lib().orc_merkle_prove(_ptr(digests), C.c_int(log_leaves), C.c_int(cap_height), C.c_size_t(leaf_index),
                                _ptr(sib))
    if rc != 0:
        raise ValueError("bad leaf index")
    return sib.reshape(-1, 4)


def merkle_verify(leaf, leaf_index, siblings, cap):
    leaf, siblings, cap = _a(leaf), _a(siblings), _a(cap)
    cap_height = int(np.log2(cap.size // 4))
    return bool(lib().orc_merkle_verify(_ptr(leaf), C.c_size_t(leaf.size), C.c_size_t(leaf_index), _ptr(siblings),
                                        C.c_int(siblings.size // 4), _ptr(cap), C.c_int(cap_height)))


def merkle_digests_to_plonky2(digests, log_leaves, cap_height):
    digests = _a(digests)
    out = np.zeros_like(digests)
    if digests.size:
        lib().orc_merkle_digests_to_plonky2(_ptr(digests), C.c_int(log_leaves), C.c_int(cap_height), _ptr(out))
    return out


def batch_from_values(values, rate_bits, cap_height, salt=None, want=("coeffs", "leaves", "digests"), from_coeffs=False):
    """values: (n_cols, n) column-major witness.  Returns dict(coeffs, leaves, digests, cap)."""
    values = _a(values)
    n_cols, n = values.shape
    log_n = int(np.log2(n))
    assert 1 << log_n == n
    nb = n << rate_bits
    n_salt = 0 if salt is None else salt.shape[0]
    salt = None if salt is None else _a(salt)
    out = {}
    coeffs = np.empty_like(values) if ("coeffs" in want and not from_coeffs) else None
    leaves = np.empty((nb, n_cols + n_salt), dtype=np.uint64) if "leaves" in want else None
    nd = n_digest_words(log_n + rate_bits, cap_height)
    digests = np.zeros(nd, dtype=np.uint64) if ("digests" in want and nd > 0) else None
    cap = np.zeros(4 << cap_height, dtype=np.uint64)
    L = lib()
    if from_coeffs:
        rc = L.orc_batch_from_coeffs(_ptr(values), C.c_int(log_n), C.c_int(n_cols), C.c_int(rate_bits),
                                     C.c_int(cap_height), _ptr(salt), C.c_int(n_salt), _ptr(leaves), _ptr(digests),
                                     _ptr(cap))
    else:
        rc = L.orc_batch_from_values(_ptr(values), C.c_int(log_n), C.c_int(n_cols), C.c_int(rate_bits),
                                     C.c_int(cap_height), _ptr(salt), C.c_int(n_salt), _ptr(coeffs), _ptr(leaves),
                                     _ptr(digests), _ptr(cap))
    if rc != 0:
        raise ValueError("orc_batch_from_values failed rc=%d" % rc)
    out["coeffs"] = values if from_coeffs else coeffs
    out["leaves"] = leaves
    out["digests"] = None if digests is None else digests.reshape(-1, 4)
    out["cap"] = cap.reshape(-1, 4)
    return out


def num_threads():
    return lib().orc_num_threads()


def set_num_threads(n):
    lib().orc_set_num_threads(C.c_int(n))


# ------------------------------------------------------------------ challenger / FRI
class Challenger(C.Structure):
    """plonky2 iop/challenger.rs `Challenger<GoldilocksField, PoseidonHash>` (oracle side)."""
    _fields_ = [("state", C.c_uint64 * 12), ("inbuf", C.c_uint64 * 8), ("outbuf", C.c_uint64 * 8),
                ("n_in", C.c_int32), ("n_out", C.c_int32)]

    def __init__(self):
        super().__init__()
        lib().orc_challenger_init(C.byref(self))

    def observe(self, xs):
        xs = _a(np.atleast_1d(xs)).reshape(-1)
        lib().orc_challenger_observe_many(C.byref(self), _ptr(xs), C.c_size_t(xs.size))

    def get_challenge(self):
        f = lib().orc_challenger_get
        f.restype = C.c_uint64
        return int(f(C.byref(self)))

    def get_extension_challenge(self):
        out = np.zeros(2, dtype=np.uint64)
        lib().orc_challenger_get_ext(C.byref(self), _ptr(out))
        return out

    def copy(self):
        c = Challenger()
        C.memmove(C.byref(c), C.byref(self), C.sizeof(Challenger))
        return c

    def snapshot(self):
        return (list(self.state), list(self.inbuf)[:self.n_in], list(self.outbuf)[:self.n_out])


class _BatchView(C.Structure):
    _fields_ = [("log_n", C.c_int32), ("n_cols", C.c_int32), ("n_salt", C.c_int32), ("rate_bits", C.c_int32),
                ("cap_height", C.c_int32), ("coeffs", C.c_void_p), ("leaves", C.c_void_p), ("digests", C.c_void_p),
                ("cap", C.c_void_p)]


class _FriInstance(C.Structure):
    _fields_ = [("n_batches", C.c_int32), ("points", C.c_void_p), ("n_polys", C.c_void_p),
                ("poly_oracle", C.c_void_p), ("poly_index", C.c_void_p)]


class _FriParams(C.Structure):
    _fields_ = [("degree_bits", C.c_int32), ("rate_bits", C.c_int32), ("cap_height", C.c_int32),
                ("proof_of_work_bits", C.c_int32), ("num_query_rounds", C.c_int32), ("n_arity", C.c_int32),
                ("arity_bits", C.c_int32 * 16)]


def _views(batches, rate_bits, cap_height):
    """batches: list of dicts from batch_from_values (coeffs, leaves, digests, cap)."""
    arr = (_BatchView * len(batches))()
    keep = []
    for i, b in enumerate(batches):
        n_cols, n = b["coeffs"].shape
        co, lv, cp = _a(b["coeffs"]), _a(b["leaves"]), _a(b["cap"])
        dg = _a(b["digests"]) if b["digests"] is not None else np.zeros(4, np.uint64)
        keep += [co, lv, cp, dg]
        arr[i] = _BatchView(int(np.log2(n)), n_cols, lv.shape[1] - n_cols, rate_bits, cap_height,
                            co.ctypes.data, lv.ctypes.data, dg.ctypes.data, cp.ctypes.data)
    return arr, keep


def _instance(points, polys):
    """points: (n_batches, 2); polys: per batch list of (oracle_index, poly_index)."""
    pts = _a(points).reshape(-1, 2)
    n_polys = np.array([len(p) for p in polys], dtype=np.int32)
    po = np.array([o for p in polys for o, _ in p], dtype=np.int32)
    pi = np.array([i for p in polys for _, i in p], dtype=np.int32)
    inst = _FriInstance(len(polys), pts.ctypes.data, n_polys.ctypes.data, po.ctypes.data, pi.ctypes.data)
    return inst, [pts, n_polys, po, pi]


def _params(degree_bits, rate_bits, cap_height, pow_bits, num_queries, arity_bits):
    p = _FriParams(degree_bits, rate_bits, cap_height, pow_bits, num_queries, len(arity_bits))
    for i, a in enumerate(arity_bits):
        p.arity_bits[i] = a
    return p


def fri_final_poly(batches, rate_bits, cap_height, points, polys, alpha):
    v, k1 = _views(batches, rate_bits, cap_height)
    inst, k2 = _instance(points, polys)
    n = batches[0]["coeffs"].shape[1]
    out = np.zeros((n, 2), dtype=np.uint64)
    al = _a(alpha)
    rc = lib().orc_fri_final_poly(v, len(batches), C.byref(inst), _ptr(al), _ptr(out))
    assert rc == 0
    return out


def fri_prove_openings(batches, points, polys, challenger, degree_bits, rate_bits, cap_height, pow_bits,
                       num_queries, arity_bits):
    """PolynomialBatch::prove_openings restated; returns the flat proof (layout in plonky2_fri_oracle.c)."""
    v, k1 = _views(batches, rate_bits, cap_height)
    inst, k2 = _instance(points, polys)
    fp = _params(degree_bits, rate_bits, cap_height, pow_bits, num_queries, arity_bits)
    f = lib().orc_fri_proof_words
    f.restype = C.c_size_t
    proof = np.zeros(f(v, len(batches), C.byref(fp)), dtype=np.uint64)
    rc = lib().orc_fri_prove_openings(v, len(batches), C.byref(inst), C.byref(challenger), C.byref(fp), _ptr(proof))
    if rc != 0:
        raise ValueError("orc_fri_prove_openings failed")
    return proof


def eval_poly_ext(batch, poly_index, z, rate_bits=0, cap_height=0):
    v, k = _views([batch], rate_bits, cap_height)
    out = np.zeros(2, dtype=np.uint64)
    zz = _a(z)
    lib().orc_eval_poly_ext(v, C.c_int(poly_index), _ptr(zz), _ptr(out))
    return out


def fri_verify(batches, points, polys, openings, challenger, degree_bits, rate_bits, cap_height, pow_bits,
               num_queries, arity_bits, proof):
    v, k1 = _views(batches, rate_bits, cap_height)
    inst, k2 = _instance(points, polys)
    fp = _params(degree_bits, rate_bits, cap_height, pow_bits, num_queries, arity_bits)
    op, pr = _a(openings), _a(proof)
    return bool(lib().orc_fri_verify(v, len(batches), C.byref(inst), _ptr(op), C.byref(challenger), C.byref(fp),
                                     _ptr(pr)))


# ---- quotient evaluation and proof bytes (plonky2_plonk_oracle.c) ----
GATE_NOOP, GATE_CONSTANT, GATE_PUBLIC_INPUT, GATE_ARITHMETIC = 0, 1, 2, 3


class _Gate(C.Structure):
    _fields_ = [("kind", C.c_int32), ("param", C.c_int32), ("row", C.c_int32), ("selector_index", C.c_int32),
                ("group_start", C.c_int32), ("group_end", C.c_int32)]


class _Circuit(C.Structure):
    _fields_ = [("degree_bits", C.c_int32), ("rate_bits", C.c_int32), ("quotient_degree_factor", C.c_int32),
                ("num_challenges", C.c_int32), ("num_wires", C.c_int32), ("num_routed_wires", C.c_int32),
                ("num_constants", C.c_int32), ("num_selectors", C.c_int32), ("num_partial_products", C.c_int32),
                ("num_gate_constraints", C.c_int32), ("n_gates", C.c_int32), ("gates", C.c_void_p), ("k_is", C.c_void_p)]


def _circuit(circuit):
    """circuit: dict with the `orc_circuit` fields; gates = list of (kind, param, row, selector_index, group_start,
    group_end); k_is = per routed wire."""
    gates = (_Gate * max(len(circuit["gates"]), 1))(*[_Gate(*g) for g in circuit["gates"]])
    k_is = _a(circuit["k_is"])
    c = _Circuit(circuit["degree_bits"], circuit["rate_bits"], circuit["quotient_degree_factor"],
                 circuit["num_challenges"], circuit["num_wires"], circuit["num_routed_wires"], circuit["num_constants"],
                 circuit["num_selectors"], circuit["num_partial_products"], circuit["num_gate_constraints"],
                 len(circuit["gates"]), C.cast(gates, C.c_void_p), k_is.ctypes.data)
    return c, [gates, k_is]


def poseidon_gate_fill(rows):
    """gates/poseidon.rs PoseidonGenerator on an (n, 135) array: inputs (wires 0..11) and swap (wire 24) are read, the
    deltas, S-box input wires and outputs are written in place."""
    assert rows.dtype == np.uint64 and rows.ndim == 2 and rows.shape[1] == 135 and rows.flags.c_contiguous
    lib().orc_poseidon_gate_fill.restype = None
    lib().orc_poseidon_gate_fill(_ptr(rows), C.c_size_t(rows.shape[0]))
    return rows


def quotient_values(circuit, constants_sigmas, wires, zs_partial_products, public_inputs_hash, betas, gammas, alphas,
                    cap_height=0):
    """compute_quotient_polys before its `coset_ifft`: (num_challenges, 2^(degree_bits + q)) values on the coset.
    The three batches are dicts from batch_from_values (with leaves)."""
    c, keep = _circuit(circuit)
    v, k1 = _views([constants_sigmas, wires, zs_partial_products], circuit["rate_bits"], cap_height)
    qdb = int(np.ceil(np.log2(circuit["quotient_degree_factor"])))
    out = np.zeros((circuit["num_challenges"], 1 << (circuit["degree_bits"] + qdb)), dtype=np.uint64)
    pih, b, g, al = _a(public_inputs_hash), _a(betas), _a(gammas), _a(alphas)
    vp = C.cast(v, C.c_void_p).value
    sz = C.sizeof(_BatchView)
    rc = lib().orc_quotient_values(C.byref(c), C.c_void_p(vp), C.c_void_p(vp + sz), C.c_void_p(vp + 2 * sz), _ptr(pih),
                                   _ptr(b), _ptr(g), _ptr(al), _ptr(out))
    if rc != 0:
        raise ValueError("orc_quotient_values failed rc=%d" % rc)
    return out


def vanishing_at(circuit, x, constants_sigmas_row, wires_row, zs_pp_row, next_zs_row, public_inputs_hash, betas, gammas,
                 alphas):
    """The reduced vanishing polynomial(s) at base-field point x from the committed polynomials' values at x (and the
    Zs at g x): what plonky2's verifier computes from the opening set."""
    c, keep = _circuit(circuit)
    rows = [_a(r) for r in (constants_sigmas_row, wires_row, zs_pp_row, next_zs_row)]
    pih, b, g, al = _a(public_inputs_hash), _a(betas), _a(gammas), _a(alphas)
    out = np.zeros(circuit["num_challenges"], dtype=np.uint64)
    rc = lib().orc_vanishing_at(C.byref(c), C.c_uint64(int(x)), _ptr(rows[0]), _ptr(rows[1]), _ptr(rows[2]), _ptr(rows[3]),
                                _ptr(pih), _ptr(b), _ptr(g), _ptr(al), _ptr(out))
    if rc != 0:
        raise ValueError("orc_vanishing_at failed")
    return out


def vanishing_at_ext(circuit, zeta, openings, public_inputs_hash, betas, gammas, alphas):
    """plonk/vanishing_poly.rs eval_vanishing_poly at the extension point zeta from an opening set (dict or object with
    constants, plonk_sigmas, wires, plonk_zs, plonk_zs_next, partial_products: (n, 2) arrays) -> (num_challenges, 2)."""
    c, keep = _circuit(circuit)
    get = (lambda k: openings[k]) if isinstance(openings, dict) else (lambda k: getattr(openings, k))
    ops = [_a(get(k)).reshape(-1) for k in ("constants", "plonk_sigmas", "wires", "plonk_zs", "plonk_zs_next", "partial_products")]
    z, pih, b, g, al = _a(zeta), _a(public_inputs_hash), _a(betas), _a(gammas), _a(alphas)
    out = np.zeros((circuit["num_challenges"], 2), dtype=np.uint64)
    rc = lib().orc_vanishing_at_ext(C.byref(c), _ptr(z), *[_ptr(o) for o in ops], _ptr(pih), _ptr(b), _ptr(g), _ptr(al), _ptr(out))
    if rc != 0:
        raise ValueError("orc_vanishing_at_ext failed")
    return out


def ext_mul(a, b):
    out = np.zeros(2, dtype=np.uint64)
    aa, bb = _a(a), _a(b)
    lib().orc_ext_mul(_ptr(aa), _ptr(bb), _ptr(out))
    return out


def eval_polys_base(coeffs, x):
    """coeffs (k, n) -> values at the base-field point x (Horner on python ints; small test sizes only)."""
    out = []
    for c in _a(coeffs):
        acc = 0
        for v in c[::-1]:
            acc = (acc * int(x) + int(v)) % P
        out.append(acc)
    return np.array(out, dtype=np.uint64)


def write_proof_with_public_inputs(wires_cap, zs_cap, quotient_cap, openings, fri_words, degree_bits, rate_bits,
                                   cap_height, pow_bits, num_queries, arity_bits, widths, public_inputs):
    """`ProofWithPublicInputs::to_bytes`; openings = the seven OpeningSet vectors, each (n, 2)."""
    fp = _params(degree_bits, rate_bits, cap_height, pow_bits, num_queries, arity_bits)
    ops = [_a(o).reshape(-1, 2) for o in openings]
    op_ptrs = (C.c_void_p * 7)(*[o.ctypes.data for o in ops])
    n_ops = (C.c_int * 7)(*[o.shape[0] for o in ops])
    w = np.ascontiguousarray(widths, dtype=np.int32)
    caps = [_a(wires_cap), _a(zs_cap), _a(quotient_cap)]
    words, pis = _a(fri_words), _a(public_inputs)
    f = lib().orc_write_proof_with_public_inputs
    f.restype = C.c_size_t
    args = [_ptr(caps[0]), _ptr(caps[1]), _ptr(caps[2]), op_ptrs, n_ops, _ptr(words), C.byref(fp),
            C.c_void_p(w.ctypes.data), C.c_int(len(widths)), _ptr(pis), C.c_int(pis.size)]
    n = f(*args, None)
    buf = np.zeros(n, dtype=np.uint8)
    f(*args, C.c_void_p(buf.ctypes.data))
    return buf.tobytes()
