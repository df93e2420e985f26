"""The gates beyond the four basic ones in the quotient evaluation: gates/poseidon.rs PoseidonGate,
gates/arithmetic_extension.rs ArithmeticExtensionGate<2>, gates/multiplication_extension.rs MulExtensionGate<2>,
gates/base_sum.rs BaseSumGate<2>, gates/exponentiation.rs ExponentiationGate, gates/reducing.rs ReducingGate<2>,
gates/reducing_extension.rs ReducingExtensionGate<2>, gates/random_access.rs RandomAccessGate<2>, gates/poseidon_mds.rs
PoseidonMdsGate<2> and gates/coset_interpolation.rs CosetInterpolationGate<2>, on a synthetic circuit with four selector groups.

CPU tier: the oracle's constraints are checked through what defines them -- a witness produced by the generators makes
the vanishing polynomial divisible by Z_H (the verifier identity holds at a random point), a single changed wire on a row
of each gate kind breaks it; PoseidonGate rows reproduce the permutation; and the naive partial rounds the restatement
uses give the constraint values of upstream's factored form for ARBITRARY wires.
GPU tier: bsx_poseidon_gate_witness equals the oracle's generator, bsx_quotient_values equals the oracle's quotient values
word for word with all fourteen gate kinds (satisfying and random wires), and whole device proofs are accepted by the
verifier restatement (which evaluates the gates over the extension field)."""
import os
import sys

import numpy as np
import pytest

from conftest import P, splitmix_field
import __graft_entry__ as G
from test_quotient import _challenges, _device_commitments, _identity_holds, _oracle_commitments, _pis_and_circuit

POSEIDON, ARITH_EXT, MUL_EXT, BASE_SUM = 4, 5, 6, 7
EXPONENTIATION, REDUCING, REDUCING_EXT, RANDOM_ACCESS, POSEIDON_MDS, COSET_INTERPOLATION = 8, 9, 10, 11, 12, 13


@pytest.fixture(scope="module")
def pkg():
    return G.load_package()


def _row_of(sc, kind):
    """first row whose gate has this kind"""
    g = next(g for g in sc.circuit.gates if g.kind == kind)
    return int(np.nonzero(sc.constants[g.selector_index] == g.row)[0][0])


# ------------------------------------------------------------------------------------------ CPU tier
def test_poseidon_gate_rows_reproduce_the_permutation(oracle):
    rows = np.zeros((64, 135), dtype=np.uint64)
    rows[:, :12] = splitmix_field(5, (64, 12))
    rows[1::2, 24] = 1
    rows[0, :12] = 0
    oracle.poseidon_gate_fill(rows)
    for r in range(64):
        st = rows[r, :12].copy()
        if rows[r, 24]:
            st[:4], st[4:8] = rows[r, 4:8], rows[r, :4]
        assert (rows[r, 12:24] == oracle.poseidon(st)).all()
        delta = [(int(rows[r, 24]) * (int(rows[r, 4 + i]) - int(rows[r, i]))) % P for i in range(4)]
        assert [int(x) for x in rows[r, 25:29]] == delta
    assert int(rows[0, 12]) == 0x3c18a9786cb0b359  # plonky2's all-zeros test vector


def test_naive_partial_rounds_give_the_factored_forms_constraints_for_any_wires():
    """hash/poseidon.rs mds_partial_layer_init / mds_partial_layer_fast against constants + S-box + MDS per round, the
    S-box inputs replaced by arbitrary `wires` as PoseidonGate::eval_unfiltered does: identical lane-0 differences and
    identical final state (model of both in tools/gen_poseidon_constants.py)."""
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
    import gen_poseidon_constants as gp
    rc = gp.round_constants()
    first, init, K, V, W = gp.fast_partial_constants(rc)
    vals = [int(x) for x in splitmix_field(77, (12 + 22,))]
    state0, wp = vals[:12], vals[12:]
    s, naive = list(state0), []
    for p in range(22):
        s = [(s[i] + rc[12 * (4 + p) + i]) % P for i in range(12)]
        naive.append((s[0] - wp[p]) % P)
        s[0] = pow(wp[p], 7, P)
        s = gp.mds(s)
    t = [(state0[i] + first[i]) % P for i in range(12)]
    t = [t[0]] + gp.mat_vec(init, t[1:])
    fast = []
    for p in range(22):
        fast.append((t[0] - wp[p]) % P)
        s0 = (pow(wp[p], 7, P) + K[p]) % P
        d = (25 * s0 + sum(V[p][i] * t[1 + i] for i in range(11))) % P
        t = [d] + [(t[1 + i] + s0 * W[p][i]) % P for i in range(11)]
    assert naive == fast and s == t


@pytest.mark.parametrize("degree_bits,kw", [(6, {}), (8, {}), (6, dict(num_wires=40, num_routed_wires=24, num_ops=5))])
def test_oracle_identity_with_every_gate_kind(pkg, oracle, degree_bits, kw):
    sc = pkg.synthetic_circuit(degree_bits, seed=13, gate_set="all", poseidon_fill=oracle.poseidon_gate_fill, **kw)
    c = sc.circuit
    kinds = sorted(g.kind for g in c.gates)
    big = c.num_wires >= 135
    assert kinds == (list(range(14)) if big else [0, 1, 2, 3, 5, 6, 7, 8, 9, 10, 11, 13])  # 40 wires: no Poseidon / PoseidonMds
    assert c.num_selectors == 4 and c.num_gate_constraints == (123 if big else 33)
    for g in c.gates:  # a run of rows of every kind
        assert (sc.constants[g.selector_index] == g.row).any(), g
    betas, gammas, alphas = _challenges(31, c.num_challenges)
    assert _identity_holds(oracle, sc, betas, gammas, alphas) == (True, True)
    # one wire of one row of each kind: (kind, wire); the Poseidon ones hit a delta, a partial-round and a last-round S-box wire
    ra = next(g.param for g in c.gates if g.kind == RANDOM_ACCESS)
    ra_bit0 = (2 + (1 << (ra & 0xFF))) * ((ra >> 8) & 0xFF) + (ra >> 16)
    n_exp = next(g.param for g in c.gates if g.kind == EXPONENTIATION)
    tampers = [(BASE_SUM, 0), (BASE_SUM, 5), (ARITH_EXT, 6), (ARITH_EXT, 1), (MUL_EXT, 5), (MUL_EXT, 2),
               (EXPONENTIATION, 0), (EXPONENTIATION, 1 + n_exp), (EXPONENTIATION, 2 + n_exp + 3), (REDUCING, 2), (REDUCING, 7),
               (REDUCING, 0), (REDUCING_EXT, 5), (REDUCING_EXT, 9), (RANDOM_ACCESS, 0), (RANDOM_ACCESS, 1),
               (RANDOM_ACCESS, ra_bit0 + 1), (RANDOM_ACCESS, ra_bit0 - 1)]
    ci = next(g.param for g in c.gates if g.kind == COSET_INTERPOLATION)
    ci_np = 1 << (ci & 0xFF)
    # shift, a value, the evaluation point, the evaluation value, the shifted point (last two wires of the gate)
    ci_end = 1 + 2 * ci_np + 4 + 4 * ((ci_np - 2) // ((ci >> 8) - 1)) + 2
    tampers += [(COSET_INTERPOLATION, 0), (COSET_INTERPOLATION, 4), (COSET_INTERPOLATION, 1 + 2 * ci_np),
                (COSET_INTERPOLATION, 3 + 2 * ci_np), (COSET_INTERPOLATION, ci_end - 1)]
    if (ci_np - 2) // ((ci >> 8) - 1):
        tampers += [(COSET_INTERPOLATION, 5 + 2 * ci_np), (COSET_INTERPOLATION, ci_end - 3)]  # an intermediate eval / prod
    if big:
        tampers += [(POSEIDON, 3), (POSEIDON, 24), (POSEIDON, 26), (POSEIDON, 40), (POSEIDON, 70), (POSEIDON, 100), (POSEIDON, 13),
                    (POSEIDON_MDS, 5), (POSEIDON_MDS, 30)]
    for kind, wire in tampers:
        row = _row_of(sc, kind)
        keep = sc.wires[wire, row]
        sc.wires[wire, row] = (int(keep) + 1) % P
        assert not _identity_holds(oracle, sc, betas, gammas, alphas)[0], (kind, wire)
        sc.wires[wire, row] = keep
    assert _identity_holds(oracle, sc, betas, gammas, alphas)[0]


def test_coset_interpolation_rows_hold_the_interpolant(pkg, oracle):
    """what the gate is for: evaluation_value = the polynomial through (shift * g^i, value_i), i < 2^bits, at evaluation_point,
    in F_p[X] / (X^2 - 7) -- plain Lagrange interpolation on python ints against the generated row"""
    sc = pkg.synthetic_circuit(6, seed=6, gate_set="all", poseidon_fill=oracle.poseidon_gate_fill)
    g = next(g for g in sc.circuit.gates if g.kind == COSET_INTERPOLATION)
    bits, npts = g.param & 0xFF, 1 << (g.param & 0xFF)
    row = _row_of(sc, COSET_INTERPOLATION)
    w = [int(x) for x in sc.wires[:, row]]
    mul = lambda a, b: ((a[0] * b[0] + 7 * a[1] * b[1]) % P, (a[0] * b[1] + a[1] * b[0]) % P)
    gen = pow(1753635133440165772, 1 << (32 - bits), P)
    xs = [w[0] * pow(gen, i, P) % P for i in range(npts)]
    z = (w[1 + 2 * npts], w[2 + 2 * npts])
    acc = (0, 0)
    for i in range(npts):
        term = (w[1 + 2 * i], w[2 + 2 * i])
        for j in range(npts):
            if j != i:
                inv = pow((xs[i] - xs[j]) % P, P - 2, P)
                term = mul(term, ((z[0] - xs[j]) * inv % P, z[1] * inv % P))
        acc = ((acc[0] + term[0]) % P, (acc[1] + term[1]) % P)
    assert acc == (w[3 + 2 * npts], w[4 + 2 * npts])


def test_base_sum_limbs_must_be_bits(pkg, oracle):
    """a limb of 2 with the sum adjusted passes the sum constraint and fails limb (limb - 1)"""
    sc = pkg.synthetic_circuit(6, seed=2, gate_set="all", poseidon_fill=oracle.poseidon_gate_fill)
    betas, gammas, alphas = _challenges(3, 2)
    row = _row_of(sc, BASE_SUM)
    old = int(sc.wires[3, row])
    sc.wires[3, row] = 2
    sc.wires[0, row] = (int(sc.wires[0, row]) + (2 - old) * 4) % P
    assert not _identity_holds(oracle, sc, betas, gammas, alphas)[0]


def test_random_access_bits_must_match_the_index(pkg, oracle):
    """the claimed element of another index with the bits left alone fails; with index and bits changed together it passes"""
    sc = pkg.synthetic_circuit(6, seed=4, gate_set="all", poseidon_fill=oracle.poseidon_gate_fill)
    betas, gammas, alphas = _challenges(4, 2)
    g = next(g for g in sc.circuit.gates if g.kind == RANDOM_ACCESS)
    bits, copies, extra = g.param & 0xFF, (g.param >> 8) & 0xFF, g.param >> 16
    row, vec = _row_of(sc, RANDOM_ACCESS), 1 << bits
    idx = int(sc.wires[0, row])
    other = (idx + 1) % vec
    sc.wires[1, row] = sc.wires[2 + other, row]
    assert not _identity_holds(oracle, sc, betas, gammas, alphas)[0]
    sc.wires[0, row] = other
    b0 = (2 + vec) * copies + extra
    sc.wires[b0:b0 + bits, row] = [(other >> i) & 1 for i in range(bits)]
    assert _identity_holds(oracle, sc, betas, gammas, alphas)[0]


def test_oracle_proof_with_every_gate_kind_verifies(pkg, oracle):
    """The whole prove() flow on the CPU oracle for a circuit with all fourteen kinds, checked by the verifier restatement:
    this is the CPU-tier test of the F_p^2 gate evaluators (ExtensionAlgebra for the extension gates), which otherwise only
    the GPU tier's whole-proof tests reach.  One changed wire on a row of each kind and the vanishing identity fails."""
    import plonk_verifier
    from test_quotient import _oracle_prove
    public_inputs, sc = _pis_and_circuit(pkg, oracle, 6, 9, gate_set="all", poseidon_fill=oracle.poseidon_gate_fill)
    cfg = pkg.FriConfig(rate_bits=sc.circuit.rate_bits, cap_height=1, proof_of_work_bits=4, num_query_rounds=3)
    pw, aux = _oracle_prove(pkg, oracle, sc, cfg, public_inputs)
    ok, why = plonk_verifier.verify(oracle, sc.circuit.as_dict(), pw, aux)
    assert ok, why
    for kind, wire in ((POSEIDON, 66), (POSEIDON_MDS, 40), (ARITH_EXT, 3), (MUL_EXT, 4), (BASE_SUM, 1), (EXPONENTIATION, 2),
                       (REDUCING, 4), (REDUCING_EXT, 6), (RANDOM_ACCESS, 1), (COSET_INTERPOLATION, 2)):
        row = _row_of(sc, kind)
        keep = sc.wires[wire, row]
        sc.wires[wire, row] = (int(keep) + 1) % P
        pw_bad, aux_bad = _oracle_prove(pkg, oracle, sc, cfg, public_inputs)
        ok, why = plonk_verifier.verify(oracle, sc.circuit.as_dict(), pw_bad, aux_bad)
        assert not ok and "vanishing" in why, (kind, wire)
        sc.wires[wire, row] = keep


# ------------------------------------------------------------------------------------------ GPU tier
@pytest.fixture(scope="module")
def ctx(pkg):
    c = pkg.Context(0)
    yield c
    c.close()


@pytest.mark.gpu
def test_device_poseidon_gate_witness_equals_the_oracle(ctx, oracle):
    rows = splitmix_field(9, (777, 135))
    rows[:, 24] = splitmix_field(10, (777,)) & np.uint64(1)
    rows[0, :12] = 0
    rows[1, :12] = P - 1
    want = oracle.poseidon_gate_fill(rows.copy())
    got = ctx.poseidon_gate_witness(rows)
    assert (got == want).all()
    assert (got[:, :12] == rows[:, :12]).all() and (got[:, 24] == rows[:, 24]).all()


@pytest.mark.gpu
@pytest.mark.parametrize("degree_bits,kw,satisfying", [
    (6, {}, True), (8, {}, True), (8, {}, False), (11, {}, True),
    (6, dict(num_wires=40, num_routed_wires=24, num_ops=5), True), (7, dict(num_challenges=3, rate_bits=4), False)])
def test_device_quotient_values_with_every_gate_kind(pkg, ctx, oracle, degree_bits, kw, satisfying):
    sc = pkg.synthetic_circuit(degree_bits, seed=50 + degree_bits, gate_set="all", poseidon_fill=ctx.poseidon_gate_witness, **kw)
    c = sc.circuit
    if not satisfying:  # every constraint of every gate takes non-zero values
        sc.wires = splitmix_field(78, sc.wires.shape)
    betas, gammas, alphas = _challenges(70 + degree_bits, c.num_challenges)
    ob = _oracle_commitments(oracle, sc, betas, gammas)
    gb = _device_commitments(pkg, ctx, sc, betas, gammas)
    want = oracle.quotient_values(c.as_dict(), ob[0], ob[1], ob[2], sc.public_inputs_hash, betas, gammas, alphas)
    got = pkg.quotient_values(ctx, c, gb[0], gb[1], gb[2], sc.public_inputs_hash, betas, gammas, alphas)
    assert got.shape == want.shape and (got == want).all()
    qb = pkg.compute_quotient_polys(ctx, c, gb[0], gb[1], gb[2], sc.public_inputs_hash, betas, gammas, alphas, 1)
    chunks = oracle.quotient_chunks(want, degree_bits, c.quotient_degree_factor)
    assert (qb.polynomials == chunks).all()
    qb.close()
    for b in gb:
        b.close()


@pytest.mark.gpu
def test_device_quotient_over_drawn_gate_parameters(pkg, ctx, oracle):
    """`hypothesis` draws the wire counts, from which the gates take their parameters (operations per extension gate,
    power bits, coefficients, limbs, the RandomAccess bits / copies / extra constants; PoseidonMds and Poseidon drop out
    below 48 / 135 wires), the challenge count and whether the witness satisfies the circuit."""
    from hypothesis import HealthCheck, given, settings, strategies as st

    @settings(max_examples=25, deadline=None, suppress_health_check=list(HealthCheck), derandomize=True)
    @given(routed=st.integers(8, 80), extra_wires=st.integers(0, 60), ops=st.integers(1, 20), num_challenges=st.integers(1, 4),
           satisfying=st.booleans(), seed=st.integers(0, 2**20))
    def run(routed, extra_wires, ops, num_challenges, satisfying, seed):
        sc = pkg.synthetic_circuit(6, seed=seed, gate_set="all", poseidon_fill=ctx.poseidon_gate_witness, num_wires=routed + extra_wires,
                                   num_routed_wires=routed, num_ops=max(1, min(routed // 4, ops)), num_challenges=num_challenges)
        c = sc.circuit
        if not satisfying:
            sc.wires = splitmix_field(seed + 1, sc.wires.shape)
        betas, gammas, alphas = _challenges(seed, num_challenges)
        ob = _oracle_commitments(oracle, sc, betas, gammas)
        gb = _device_commitments(pkg, ctx, sc, betas, gammas)
        try:
            want = oracle.quotient_values(c.as_dict(), ob[0], ob[1], ob[2], sc.public_inputs_hash, betas, gammas, alphas)
            got = pkg.quotient_values(ctx, c, gb[0], gb[1], gb[2], sc.public_inputs_hash, betas, gammas, alphas)
            assert got.shape == want.shape and (got == want).all(), [(g.kind, g.param) for g in c.gates]
        finally:
            for b in gb:
                b.close()

    run()


@pytest.mark.gpu
def test_poseidon_gate_needs_135_wires(pkg, ctx):
    sc = pkg.synthetic_circuit(5, seed=2, num_wires=40, num_routed_wires=24, num_ops=5)
    c = sc.circuit
    betas, gammas, alphas = _challenges(5, 2)
    gb = _device_commitments(pkg, ctx, sc, betas, gammas)
    c.gates[0] = pkg.Gate(pkg.GATE_POSEIDON, 0, 0, 0, 0, 4)
    with pytest.raises(pkg.BsxError, match="does not fit"):
        pkg.quotient_values(ctx, c, gb[0], gb[1], gb[2], sc.public_inputs_hash, betas, gammas, alphas)
    for b in gb:
        b.close()


@pytest.mark.gpu
@pytest.mark.parametrize("degree_bits", [6, 10])
def test_device_proof_with_every_gate_kind_is_accepted(pkg, ctx, oracle, degree_bits):
    """prove() on the device for a circuit with Poseidon / extension / base-sum rows: the verifier restatement (gates
    evaluated over F_p^2 at zeta) accepts it; one changed S-box wire of one PoseidonGate row and it does not."""
    import plonk_verifier
    public_inputs, sc = _pis_and_circuit(pkg, oracle, degree_bits, 7, gate_set="all", poseidon_fill=ctx.poseidon_gate_witness)
    cfg = pkg.FriConfig(rate_bits=sc.circuit.rate_bits, cap_height=2, proof_of_work_bits=8, num_query_rounds=6)
    pw, aux = pkg.prove(ctx, sc, cfg, public_inputs)
    ok, why = plonk_verifier.verify(oracle, sc.circuit.as_dict(), pw, aux)
    assert ok, why
    for kind, wire in ((POSEIDON, 77), (ARITH_EXT, 7), (BASE_SUM, 2), (EXPONENTIATION, 0), (REDUCING_EXT, 3), (RANDOM_ACCESS, 1),
                       (POSEIDON_MDS, 47), (COSET_INTERPOLATION, 9)):
        row = _row_of(sc, kind)
        keep = sc.wires[wire, row]
        sc.wires[wire, row] = (int(keep) + 1) % P
        pw_bad, aux_bad = pkg.prove(ctx, sc, cfg, public_inputs)
        ok, why = plonk_verifier.verify(oracle, sc.circuit.as_dict(), pw_bad, aux_bad)
        assert not ok and "vanishing" in why, (kind, wire)
        sc.wires[wire, row] = keep
