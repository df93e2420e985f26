"""Quotient evaluation: plonk/prover.rs compute_quotient_polys + plonk/vanishing_poly.rs
eval_vanishing_poly_base_batch on a synthetic circuit of the four built-in gates.

CPU tier: the oracle's restatement is checked through the property that defines it -- for a witness that satisfies
every gate and copy constraint the vanishing polynomial is divisible by Z_H, i.e. the quotient's coefficients beyond
quotient_degree_factor * N vanish (`trim_to_len` succeeds), and for a tampered witness they do not.
GPU tier: the device evaluator equals the oracle word for word, and the device pipeline (values -> coset_ifft -> trim
-> chunks -> commitment) ends in the oracle's quotient commitment."""
import numpy as np
import pytest

from conftest import P, splitmix_field
import __graft_entry__ as G


@pytest.fixture(scope="module")
def pkg():
    return G.load_package()


def _oracle_commitments(oracle, sc, betas, gammas, cap_height=1):
    c = sc.circuit
    cs = oracle.batch_from_values(np.concatenate([sc.constants, sc.sigmas]), c.rate_bits, cap_height)
    wires = oracle.batch_from_values(sc.wires, c.rate_bits, cap_height)
    zs_vals = oracle.partial_products_and_zs(sc.wires[:c.num_routed_wires], sc.sigmas, c.k_is, betas, gammas,
                                             c.quotient_degree_factor)
    zs = oracle.batch_from_values(zs_vals, c.rate_bits, cap_height)
    return cs, wires, zs, zs_vals


def _challenges(seed, nc):
    v = splitmix_field(seed, (3, nc))
    return v[0], v[1], v[2]


def _identity_holds(oracle, sc, betas, gammas, alphas, seed=0):
    """plonky2's verifier identity at a random base-field point zeta, from the oracle's own prover output:
    V(zeta) == Z_H(zeta) * sum_i zeta^(n i) chunk_i(zeta), V from the polynomials' values at zeta (and Z at g zeta).
    With quotient_degree_factor = 2^q `trim_to_len` checks nothing, so this is the test that the quotient values are
    those of a polynomial multiple of Z_H.  Returns (holds, trim_ok)."""
    c = sc.circuit
    n, nc, f = 1 << c.degree_bits, c.num_challenges, c.quotient_degree_factor
    cs, wires, zs, _ = _oracle_commitments(oracle, sc, betas, gammas)
    q = oracle.quotient_values(c.as_dict(), cs, wires, zs, sc.public_inputs_hash, betas, gammas, alphas)
    coeffs = np.stack([oracle.coset_ifft(v) for v in q])
    trim_ok = not coeffs[:, f * n:].any()
    chunks = coeffs[:, :f * n].reshape(nc * f, n)
    zeta = int(splitmix_field(1234 + seed, (1,))[0])
    g = pow(1753635133440165772, 1 << (32 - c.degree_bits), P)
    rows = [oracle.eval_polys_base(b["coeffs"], zeta) for b in (cs, wires, zs)]
    next_zs = oracle.eval_polys_base(zs["coeffs"][:nc], zeta * g % P)
    v = oracle.vanishing_at(c.as_dict(), zeta, rows[0], rows[1], rows[2], next_zs, sc.public_inputs_hash, betas, gammas, alphas)
    z_h, zn = (pow(zeta, n, P) - 1) % P, pow(zeta, n, P)
    ce = [int(x) for x in oracle.eval_polys_base(chunks, zeta)]
    ok = True
    for a in range(nc):
        acc = 0
        for t in reversed(ce[a * f:(a + 1) * f]):  # reduce_with_powers(chunk evaluations, zeta^n)
            acc = (acc * zn + t) % P
        ok = ok and int(v[a]) == z_h * acc % P
    return ok, trim_ok


KW = [(5, {}), (7, {}), (6, dict(num_wires=20, num_routed_wires=12, num_ops=3, quotient_degree_factor=8)),
      (6, dict(num_wires=30, num_routed_wires=22, num_ops=5, quotient_degree_factor=5, num_challenges=3))]


@pytest.mark.parametrize("degree_bits,kw", KW)
def test_oracle_quotient_is_divisible_for_a_satisfying_witness(pkg, oracle, degree_bits, kw):
    sc = pkg.synthetic_circuit(degree_bits, seed=11, **kw)
    c = sc.circuit
    betas, gammas, alphas = _challenges(21, c.num_challenges)
    zs_vals = _oracle_commitments(oracle, sc, betas, gammas)[3]
    assert (zs_vals[:c.num_challenges, 0] == 1).all()  # Z(1) = 1
    assert (sc.sigmas != _identity_sigmas(c)).any()    # the circuit has copy constraints
    assert _identity_holds(oracle, sc, betas, gammas, alphas) == (True, True)
    partial = c.quotient_degree_factor < (1 << c.quotient_degree_bits)  # only then does trim_to_len check anything
    # break one arithmetic output: the gate constraints stop vanishing on H
    bad = pkg.synthetic_circuit(degree_bits, seed=11, **kw)
    row = int(np.nonzero(bad.constants[0] == 3)[0][0])
    bad.wires[3, row] = (int(bad.wires[3, row]) + 1) % P
    ok, trim_ok = _identity_holds(oracle, bad, betas, gammas, alphas)
    assert not ok and trim_ok == (not partial)
    # break a copy constraint only (a free input that is wired to another cell): the permutation argument fails
    bad = pkg.synthetic_circuit(degree_bits, seed=11, **kw)
    j, r = [int(v[0]) for v in np.nonzero(bad.sigmas != _identity_sigmas(c))]
    assert j % 4 != 3
    i = j // 4
    bad.wires[j, r] = (int(bad.wires[j, r]) + 5) % P
    c0, c1 = int(bad.constants[1, r]), int(bad.constants[2, r])  # keep the gate itself satisfied
    bad.wires[4 * i + 3, r] = (int(bad.wires[4 * i, r]) * int(bad.wires[4 * i + 1, r]) % P * c0 + int(bad.wires[4 * i + 2, r]) * c1) % P
    ok, trim_ok = _identity_holds(oracle, bad, betas, gammas, alphas)
    assert not ok and trim_ok == (not partial)


def _identity_sigmas(c):
    w = pow(1753635133440165772, 1 << (32 - c.degree_bits), P)
    return np.array([[int(k) * pow(w, r, P) % P for r in range(1 << c.degree_bits)] for k in c.k_is], dtype=np.uint64)


def test_public_input_and_constant_gates_bind(pkg, oracle):
    betas, gammas, alphas = _challenges(8, 2)
    for tamper in ("pi", "const"):
        bad = pkg.synthetic_circuit(5, seed=3)
        if tamper == "pi":
            bad.public_inputs_hash[2] ^= np.uint64(1)
        else:
            row = int(np.nonzero(bad.constants[0] == 1)[0][0])
            bad.wires[1, row] = (int(bad.wires[1, row]) + 1) % P
        assert _identity_holds(oracle, bad, betas, gammas, alphas)[0] is False


# ------------------------------------------------------------------------------------------ GPU tier
@pytest.fixture(scope="module")
def ctx(pkg):
    c = pkg.Context(0)
    yield c
    c.close()


def _device_commitments(pkg, ctx, sc, betas, gammas, cap_height=1):
    c = sc.circuit
    cs = pkg.PolynomialBatch.from_values(ctx, np.concatenate([sc.constants, sc.sigmas]), c.rate_bits, False, cap_height)
    wires = pkg.PolynomialBatch.from_values(ctx, sc.wires, c.rate_bits, False, cap_height)
    zs_vals = ctx.partial_products_and_zs(sc.wires[:c.num_routed_wires], sc.sigmas, c.k_is, betas, gammas,
                                          c.quotient_degree_factor)
    zs = pkg.PolynomialBatch.from_values(ctx, zs_vals, c.rate_bits, False, cap_height)
    return cs, wires, zs


@pytest.mark.gpu
@pytest.mark.parametrize("degree_bits,kw,satisfying", [
    (5, {}, True), (9, {}, True), (9, {}, False), (12, {}, True),
    (6, dict(num_wires=20, num_routed_wires=12, num_ops=3, quotient_degree_factor=8), True),
    (7, dict(num_wires=30, num_routed_wires=22, num_ops=5, quotient_degree_factor=5, num_challenges=3), True),
    (7, dict(num_wires=30, num_routed_wires=22, num_ops=5, quotient_degree_factor=4, num_challenges=1, rate_bits=2), False)])
def test_device_quotient_values_equal_the_oracle(pkg, ctx, oracle, degree_bits, kw, satisfying):
    sc = pkg.synthetic_circuit(degree_bits, seed=40 + degree_bits, **kw)
    c = sc.circuit
    if not satisfying:  # random wires: every term of the vanishing polynomial is exercised with non-zero values
        sc.wires = splitmix_field(77, sc.wires.shape)
    betas, gammas, alphas = _challenges(60 + degree_bits, c.num_challenges)
    ob = _oracle_commitments(oracle, sc, betas, gammas)
    gb = _device_commitments(pkg, ctx, sc, betas, gammas)
    for o, g in zip(ob[:3], gb):
        assert (o["cap"] == g.cap).all()
    want = oracle.quotient_values(c.as_dict(), ob[0], ob[1], ob[2], sc.public_inputs_hash, betas, gammas, alphas)
    got = pkg.quotient_values(ctx, c, gb[0], gb[1], gb[2], sc.public_inputs_hash, betas, gammas, alphas)
    assert got.shape == want.shape and (got == want).all()
    vacuous = c.quotient_degree_factor == 1 << c.quotient_degree_bits  # trim_to_len drops nothing: no check upstream either
    if satisfying or vacuous:
        qb = pkg.compute_quotient_polys(ctx, c, gb[0], gb[1], gb[2], sc.public_inputs_hash, betas, gammas, alphas, 1)
        chunks = oracle.quotient_chunks(want, degree_bits, c.quotient_degree_factor)
        assert (qb.polynomials == chunks).all()
        assert (qb.cap == oracle.batch_from_values(chunks, c.rate_bits, 1, from_coeffs=True)["cap"]).all()
        qb.close()
    else:
        with pytest.raises(pkg.BsxError, match="trim_to_len"):
            pkg.compute_quotient_polys(ctx, c, gb[0], gb[1], gb[2], sc.public_inputs_hash, betas, gammas, alphas, 1)
    for b in gb:
        b.close()


@pytest.mark.gpu
def test_quotient_rejects_mismatched_shapes(pkg, ctx):
    sc = pkg.synthetic_circuit(5, seed=2)
    c = sc.circuit
    betas, gammas, alphas = _challenges(5, 2)
    gb = _device_commitments(pkg, ctx, sc, betas, gammas)
    used = ctx.mem_used()
    with pytest.raises(pkg.BsxError, match="shape"):
        pkg.quotient_values(ctx, c, gb[1], gb[1], gb[2], sc.public_inputs_hash, betas, gammas, alphas)
    c2 = pkg.synthetic_circuit(5, seed=2).circuit
    c2.gates[3].param = 40  # 160 wires
    with pytest.raises(pkg.BsxError, match="does not fit"):
        pkg.quotient_values(ctx, c2, gb[0], gb[1], gb[2], sc.public_inputs_hash, betas, gammas, alphas)
    assert ctx.mem_used() == used
    for b in gb:
        b.close()
