"""Tier 2 for the FRI half: challenger, final polynomial, and whole FRI proofs bit-exact vs the oracle."""
import numpy as np
import pytest

from conftest import P, splitmix_field
import __graft_entry__ as G

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def pkg():
    return G.load_package()


@pytest.fixture(scope="module")
def ctx(pkg):
    c = pkg.Context(0)
    yield c
    c.close()


def test_challenger_matches_oracle(pkg, ctx, oracle):
    a, b = pkg.Challenger(ctx), oracle.Challenger()
    xs = splitmix_field(31, (29,))
    for chunk in (xs[:3], xs[3:11], xs[11:12], xs[12:29]):
        a.observe_elements(chunk)
        b.observe(chunk)
        assert a.snapshot() == b.snapshot()
        got = a.get_n_challenges(3)
        assert [int(v) for v in got] == [b.get_challenge() for _ in range(3)]
        assert a.snapshot() == b.snapshot()
    # run past the 8 buffered outputs
    assert [int(v) for v in a.get_n_challenges(11)] == [b.get_challenge() for _ in range(11)]


def _setup(pkg, ctx, oracle, degree_bits, rate_bits, cap_height, widths, coeff_oracle=False, seed=5):
    obatches, gbatches = [], []
    for i, w in enumerate(widths):
        vals = splitmix_field(seed + i, (w, 1 << degree_bits))
        from_coeffs = coeff_oracle and i == len(widths) - 1  # last oracle built like the quotient batch
        obatches.append(oracle.batch_from_values(vals, rate_bits, cap_height, from_coeffs=from_coeffs))
        mk = pkg.PolynomialBatch.from_coeffs if from_coeffs else pkg.PolynomialBatch.from_values
        gbatches.append(mk(ctx, vals, rate_bits, False, cap_height))
        assert (gbatches[-1].cap == obatches[-1]["cap"]).all()
    zeta = splitmix_field(seed + 100, (2,))
    g = oracle.lib().orc_root_of_unity(degree_bits)
    gz = np.array([oracle.lib().orc_mul(int(zeta[0]), g), oracle.lib().orc_mul(int(zeta[1]), g)], dtype=np.uint64)
    all_polys = [(o, j) for o, w in enumerate(widths) for j in range(w)]
    polys = [all_polys, [(0, j) for j in range(min(2, widths[0]))]]
    points = np.stack([zeta, gz])
    return obatches, gbatches, points, polys


@pytest.mark.parametrize("degree_bits,widths,coeff_oracle", [(5, [2, 2], False), (10, [5, 3, 2], True), (13, [9, 4], True)])
def test_final_poly_vs_oracle(pkg, ctx, oracle, degree_bits, widths, coeff_oracle):
    ob, gb, points, polys = _setup(pkg, ctx, oracle, degree_bits, 1, 0, widths, coeff_oracle)
    alpha = splitmix_field(9, (2,))
    want = oracle.fri_final_poly(ob, 1, 0, points, polys, alpha)
    got = pkg.final_poly(pkg.FriInstanceInfo(points, polys), gb, alpha)
    assert (got == want).all()


@pytest.mark.parametrize("degree_bits,rate_bits,cap_height,pow_bits,queries,widths", [
    (5, 2, 0, 3, 4, [2, 1]),       # no reduction rounds: the proof is final poly + openings
    (6, 3, 1, 4, 5, [3, 2]),
    (9, 3, 4, 8, 7, [6, 2, 2]),
    (12, 3, 4, 10, 28, [20, 4, 3]),  # standard_recursion_config shape at a small degree
    (14, 1, 4, 6, 9, [3, 2]),
])
def test_prove_openings_bit_exact(pkg, ctx, oracle, degree_bits, rate_bits, cap_height, pow_bits, queries, widths):
    ob, gb, points, polys = _setup(pkg, ctx, oracle, degree_bits, rate_bits, cap_height, widths, coeff_oracle=True)
    cfg = pkg.FriConfig(rate_bits=rate_bits, cap_height=cap_height, proof_of_work_bits=pow_bits,
                        num_query_rounds=queries)
    fp = cfg.fri_params(degree_bits)
    och, gch = oracle.Challenger(), pkg.Challenger(ctx)
    for o, g in zip(ob, gb):
        och.observe(o["cap"])
        gch.observe_cap(g.cap)
    och_v = och.copy()
    want = oracle.fri_prove_openings(ob, points, polys, och, degree_bits, rate_bits, cap_height, pow_bits, queries,
                                     fp.reduction_arity_bits)
    got = pkg.prove_openings(pkg.FriInstanceInfo(points, polys), gb, gch, fp)
    assert got.size == want.size == pkg.fri_proof_words(gb, fp)
    assert (got == want).all()
    assert gch.snapshot() == och.snapshot()  # transcript continues identically
    # and the verifier restatement accepts the device proof
    openings = np.array([oracle.eval_poly_ext(ob[o], j, points[b]) for b, pl in enumerate(polys) for (o, j) in pl],
                        dtype=np.uint64)
    assert oracle.fri_verify(ob, points, polys, openings, och_v, degree_bits, rate_bits, cap_height, pow_bits,
                             queries, fp.reduction_arity_bits, got)


def test_fri_proof_entry_point(pkg, ctx, oracle):
    """fri::prover::fri_proof with caller-supplied LDE coefficients equals prove_openings' tail."""
    degree_bits, rate_bits, cap_height = 8, 3, 2
    ob, gb, points, polys = _setup(pkg, ctx, oracle, degree_bits, rate_bits, cap_height, [3, 2])
    cfg = pkg.FriConfig(rate_bits=rate_bits, cap_height=cap_height, proof_of_work_bits=5, num_query_rounds=6)
    fp = cfg.fri_params(degree_bits)
    gch = pkg.Challenger(ctx)
    gch.observe_cap(gb[0].cap)
    ref_ch = gch.copy()
    want = pkg.prove_openings(pkg.FriInstanceInfo(points, polys), gb, ref_ch, fp)
    alpha = gch.get_extension_challenge()
    fin = pkg.final_poly(pkg.FriInstanceInfo(points, polys), gb, alpha)
    lde_coeffs = np.zeros((1 << (degree_bits + rate_bits), 2), dtype=np.uint64)
    lde_coeffs[:1 << degree_bits] = fin
    got = pkg.fri_proof(gb, lde_coeffs, gch, fp)
    assert (got == want).all() and gch.snapshot() == ref_ch.snapshot()


def test_fri_errors(pkg, ctx):
    vals = splitmix_field(1, (2, 64))
    b = pkg.PolynomialBatch.from_values(ctx, vals, 1, False, 0)
    fp = pkg.FriConfig(rate_bits=2, cap_height=0).fri_params(6)  # rate mismatch with the oracle
    with pytest.raises(pkg.BsxError, match="does not match"):
        pkg.fri_proof_words([b], fp)
    inst = pkg.FriInstanceInfo(np.zeros((1, 2), dtype=np.uint64), [[(0, 5)]])
    fp = pkg.FriConfig(rate_bits=1, cap_height=0, proof_of_work_bits=1, num_query_rounds=1).fri_params(6)
    with pytest.raises(pkg.BsxError, match="does not exist"):
        pkg.prove_openings(inst, [b], pkg.Challenger(ctx), fp)


def test_failed_calls_release_their_temporaries(pkg, ctx, oracle):
    """Error paths (upstream: panics) must not leak device memory: pool usage is unchanged after a failed call."""
    ob, gb, points, polys = _setup(pkg, ctx, oracle, 10, 1, 0, [3, 2])
    params = pkg.FriConfig(rate_bits=1, cap_height=0, proof_of_work_bits=4, num_query_rounds=3).fri_params(10)
    good = pkg.FriInstanceInfo(points, polys)
    ch = pkg.Challenger(ctx)
    pkg.prove_openings(good, gb, ch, params)  # warm: cached tables are allocated once
    used = ctx.mem_used()
    bad = pkg.FriInstanceInfo(points, [polys[0][:-1] + [(0, 99)], polys[1]])  # polynomial index out of range
    for _ in range(3):
        with pytest.raises(pkg.BsxError):
            pkg.prove_openings(bad, gb, pkg.Challenger(ctx), params)
    assert ctx.mem_used() == used
    with pytest.raises(pkg.BsxError):  # 2^24 x 60000 words cannot be allocated
        pkg.PolynomialBatch.from_values(ctx, None, 3, False, 4, device_ptr=1 << 40, shape=(60000, 1 << 24))
    assert ctx.mem_used() == used
    pkg.prove_openings(good, gb, pkg.Challenger(ctx), params)
    assert ctx.mem_used() == used
    for b in gb:
        b.close()


def test_fri_proof_bytes_from_device_proof(pkg, ctx, oracle):
    """write_fri_proof bytes of a device proof = bytes of the oracle's proof; read_fri_proof restores the words."""
    ser = __import__(pkg.__name__ + ".serialization", fromlist=["x"])
    ob, gb, points, polys = _setup(pkg, ctx, oracle, 10, 2, 1, [4, 3])
    cfg = pkg.FriConfig(rate_bits=2, cap_height=1, proof_of_work_bits=6, num_query_rounds=5)
    fp = cfg.fri_params(10)
    proof = pkg.prove_openings(pkg.FriInstanceInfo(points, polys), gb, pkg.Challenger(ctx), fp)
    want = oracle.fri_prove_openings(ob, points, polys, oracle.Challenger(), 10, 2, 1, 6, 5, list(fp.reduction_arity_bits))
    data = pkg.fri_proof_to_bytes(proof, gb, fp)
    st = ser.fri_proof_struct(want, [4, 3], 10, 2, 1, list(fp.reduction_arity_bits), 5)
    assert data == ser.write_fri_proof(st)
    assert (ser.read_fri_proof(data, [4, 3], 10, 2, 1, list(fp.reduction_arity_bits), 5) == proof).all()
    assert pkg.fri_proof_struct(proof, gb, fp)["pow_witness"] == int(want[-1])
    for b in gb:
        b.close()
