"""Quotient evaluation: plonk/prover.rs compute_quotient_polys + plonk/vanishing_poly.rs
eval_vanishing_poly_base_batch on a synthetic circuit of the four base-field gates (the other nine kinds: tests/test_gates.py).

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
def test_device_quotient_over_drawn_circuit_shapes(pkg, ctx, oracle):
    """`hypothesis` draws the circuit's shape -- rows, wires, routed wires (down to one chunk: no partial products),
    operations per ArithmeticGate, chunk size, challenges, blowup -- and whether the witness satisfies it; quotient values,
    chunks and cap must equal the oracle's for every draw."""
    from hypothesis import HealthCheck, given, settings, strategies as st

    @settings(max_examples=30, deadline=None, suppress_health_check=list(HealthCheck), derandomize=True)
    @given(degree_bits=st.integers(2, 8), factor=st.integers(5, 8), num_ops=st.integers(1, 6), extra_routed=st.integers(0, 11),
           extra_wires=st.integers(0, 8), num_challenges=st.integers(1, 4), extra_rate=st.integers(0, 1), satisfying=st.booleans(),
           seed=st.integers(0, 2**20))
    def run(degree_bits, factor, num_ops, extra_routed, extra_wires, num_challenges, extra_rate, satisfying, seed):
        routed = max(4, 4 * num_ops) + extra_routed
        rate_bits = (factor - 1).bit_length() + extra_rate
        sc = pkg.synthetic_circuit(degree_bits, seed=seed, num_wires=routed + extra_wires, num_routed_wires=routed, num_ops=num_ops,
                                   quotient_degree_factor=factor, rate_bits=rate_bits, num_challenges=num_challenges)
        c = sc.circuit
        if not satisfying:
            sc.wires = splitmix_field(seed + 1, sc.wires.shape)
        betas, gammas, alphas = _challenges(seed, num_challenges)
        ob = _oracle_commitments(oracle, sc, betas, gammas)
        gb = _device_commitments(pkg, ctx, sc, betas, gammas)
        try:
            want = oracle.quotient_values(c.as_dict(), ob[0], ob[1], ob[2], sc.public_inputs_hash, betas, gammas, alphas)
            got = pkg.quotient_values(ctx, c, gb[0], gb[1], gb[2], sc.public_inputs_hash, betas, gammas, alphas)
            assert got.shape == want.shape and (got == want).all()
            if satisfying or factor == 8:
                qb = pkg.compute_quotient_polys(ctx, c, gb[0], gb[1], gb[2], sc.public_inputs_hash, betas, gammas, alphas, 1)
                chunks = oracle.quotient_chunks(want, degree_bits, factor)
                assert (qb.polynomials == chunks).all()
                assert (qb.cap == oracle.batch_from_values(chunks, rate_bits, 1, from_coeffs=True)["cap"]).all()
                qb.close()
            else:
                with pytest.raises(pkg.BsxError, match="trim_to_len"):
                    pkg.compute_quotient_polys(ctx, c, gb[0], gb[1], gb[2], sc.public_inputs_hash, betas, gammas, alphas, 1)
        finally:
            for b in gb:
                b.close()

    run()
    used = ctx.mem_used()  # the context keeps the transform tables of every size it has seen
    run()                  # the same draws again (derandomize): nothing new to cache, so nothing may stay behind
    assert ctx.mem_used() == used


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


# ------------------------------------------------------------------------------------------ whole proofs
def _pis_and_circuit(pkg, oracle, degree_bits, seed, **kw):
    public_inputs = splitmix_field(900 + seed, (5,))
    pih = oracle.hash_no_pad(public_inputs)
    return public_inputs, pkg.synthetic_circuit(degree_bits, seed=seed, public_inputs_hash=pih, **kw)


@pytest.mark.gpu
@pytest.mark.parametrize("degree_bits,kw", [(6, {}), (10, {}), (12, {}),
                                            (8, dict(num_wires=30, num_routed_wires=22, num_ops=5, quotient_degree_factor=5, num_challenges=3))])
def test_device_proof_is_accepted_by_the_verifier_restatement(pkg, ctx, oracle, degree_bits, kw):
    """plonk/prover.rs prove() on the device for a satisfying synthetic witness -> ProofWithPublicInputs; the oracle-side
    verifier (tests/plonk_verifier.py: transcript, vanishing identity at zeta in the extension field, FRI) accepts it;
    the bytes round-trip; a proof for a broken witness is rejected at the vanishing identity."""
    import plonk_verifier
    public_inputs, sc = _pis_and_circuit(pkg, oracle, degree_bits, 7, **kw)
    cfg = pkg.FriConfig(rate_bits=sc.circuit.rate_bits, cap_height=2, proof_of_work_bits=8, num_query_rounds=6)
    assert (pkg.hash_no_pad(ctx, public_inputs) == oracle.hash_no_pad(public_inputs)).all()
    pw, aux = pkg.prove(ctx, sc, cfg, public_inputs)
    ok, why = plonk_verifier.verify(oracle, sc.circuit.as_dict(), pw, aux)
    assert ok, why
    c = sc.circuit
    n_open = [c.num_constants, c.num_routed_wires, c.num_wires, c.num_challenges, c.num_challenges,
              c.num_challenges * c.num_partial_products, c.num_challenges * c.quotient_degree_factor]
    data = pw.to_bytes(aux["fri_params"], aux["widths"])
    back = pkg.ProofWithPublicInputs.from_bytes(data, aux["fri_params"], aux["widths"], n_open)
    ok, why = plonk_verifier.verify(oracle, c.as_dict(), back, aux)
    assert ok, why
    # a wrong public input changes the transcript and the PublicInputGate's constraint
    back.public_inputs[0] ^= np.uint64(1)
    assert not plonk_verifier.verify(oracle, c.as_dict(), back, aux)[0]
    # a witness that breaks one gate: with quotient_degree_factor = 2^q the prover cannot notice (upstream's neither);
    # the verifier must
    if c.quotient_degree_factor == 1 << c.quotient_degree_bits:
        row = int(np.nonzero(sc.constants[0] == 3)[0][0])
        sc.wires[3, row] = (int(sc.wires[3, row]) + 1) % P
        pw_bad, aux_bad = pkg.prove(ctx, sc, cfg, public_inputs)
        ok, why = plonk_verifier.verify(oracle, c.as_dict(), pw_bad, aux_bad)
        assert not ok and "vanishing" in why


def _oracle_prove(pkg, oracle, sc, cfg, public_inputs):
    """The same prove() flow as pkg.prove, every step on the CPU oracle (the reference arm of the whole-proof tests)."""
    c = sc.circuit
    nc, rb, chh = c.num_challenges, cfg.rate_bits, cfg.cap_height
    fp = cfg.fri_params(c.degree_bits)
    cs = oracle.batch_from_values(np.concatenate([sc.constants, sc.sigmas]), rb, chh)
    digest = oracle.hash_no_pad(np.concatenate([cs["cap"].reshape(-1), np.array([c.degree_bits], dtype=np.uint64)]))
    pih = oracle.hash_no_pad(public_inputs)
    ch = oracle.Challenger()
    ch.observe(digest)
    ch.observe(pih)
    wires = oracle.batch_from_values(sc.wires, rb, chh)
    ch.observe(wires["cap"].reshape(-1))
    betas = [ch.get_challenge() for _ in range(nc)]
    gammas = [ch.get_challenge() for _ in range(nc)]
    zs_vals = oracle.partial_products_and_zs(sc.wires[:c.num_routed_wires], sc.sigmas, c.k_is, betas, gammas, c.quotient_degree_factor)
    zs = oracle.batch_from_values(zs_vals, rb, chh)
    ch.observe(zs["cap"].reshape(-1))
    alphas = [ch.get_challenge() for _ in range(nc)]
    qv = oracle.quotient_values(c.as_dict(), cs, wires, zs, pih, betas, gammas, alphas, chh)
    quotient = oracle.batch_from_values(oracle.quotient_chunks(qv, c.degree_bits, c.quotient_degree_factor), rb, chh, from_coeffs=True)
    ch.observe(quotient["cap"].reshape(-1))
    zeta = ch.get_extension_challenge()
    g = pow(1753635133440165772, 1 << (32 - c.degree_bits), P)
    gz = np.array([int(zeta[0]) * g % P, int(zeta[1]) * g % P], dtype=np.uint64)
    ev = lambda b, z: np.array([oracle.eval_poly_ext(b, j, z, rb, chh) for j in range(b["coeffs"].shape[0])], dtype=np.uint64).reshape(-1, 2)
    e_cs, e_w, e_z, e_q, e_zn = ev(cs, zeta), ev(wires, zeta), ev(zs, zeta), ev(quotient, zeta), ev(zs, gz)
    openings = pkg.OpeningSet(e_cs[:c.num_constants], e_cs[c.num_constants:], e_w, e_z[:nc], e_zn[:nc], e_z[nc:], e_q)
    for batch in openings.to_fri_openings():
        ch.observe(batch.reshape(-1))
    batches = [cs, wires, zs, quotient]
    widths = [b["coeffs"].shape[0] for b in batches]
    polys = [[(o, j) for o, w in enumerate(widths) for j in range(w)], [(2, j) for j in range(nc)]]
    words = oracle.fri_prove_openings(batches, np.stack([zeta, gz]), polys, ch, c.degree_bits, rb, chh, cfg.proof_of_work_bits,
                                      cfg.num_query_rounds, list(fp.reduction_arity_bits))
    pw = pkg.ProofWithPublicInputs(pkg.Proof(wires["cap"], zs["cap"], quotient["cap"], openings, words), np.asarray(public_inputs, dtype=np.uint64))
    return pw, {"circuit_digest": digest, "constants_sigmas_cap": cs["cap"], "fri_params": fp, "widths": widths}


@pytest.mark.parametrize("degree_bits,kw", [(5, {}), (6, dict(num_wires=30, num_routed_wires=22, num_ops=5, quotient_degree_factor=5, num_challenges=3))])
def test_oracle_proof_is_accepted_by_the_verifier_restatement(pkg, oracle, degree_bits, kw):
    """CPU tier of the whole-proof check: prover steps and verifier are separate restatements (base-field batch evaluation
    on the coset vs extension-field evaluation of the opening set); they must agree on a satisfying witness and disagree
    on a broken one."""
    import plonk_verifier
    public_inputs, sc = _pis_and_circuit(pkg, oracle, degree_bits, 3, **kw)
    cfg = pkg.FriConfig(rate_bits=sc.circuit.rate_bits, cap_height=1, proof_of_work_bits=4, num_query_rounds=4)
    pw, aux = _oracle_prove(pkg, oracle, sc, cfg, public_inputs)
    ok, why = plonk_verifier.verify(oracle, sc.circuit.as_dict(), pw, aux)
    assert ok, why
    # an opening that is not the polynomial's value: FRI rejects; a tampered cap: the transcript diverges
    pw.proof.openings.wires[0, 0] ^= np.uint64(1)
    assert not plonk_verifier.verify(oracle, sc.circuit.as_dict(), pw, aux)[0]
    pw.proof.openings.wires[0, 0] ^= np.uint64(1)
    pw.proof.wires_cap = pw.proof.wires_cap.copy()
    pw.proof.wires_cap[0, 0] ^= np.uint64(1)
    assert not plonk_verifier.verify(oracle, sc.circuit.as_dict(), pw, aux)[0]
    if sc.circuit.quotient_degree_factor == 8:  # broken gate, unnoticed by the prover: rejected at the vanishing identity
        row = int(np.nonzero(sc.constants[0] == 3)[0][0])
        sc.wires[3, row] = (int(sc.wires[3, row]) + 1) % P
        pw_bad, aux_bad = _oracle_prove(pkg, oracle, sc, cfg, public_inputs)
        ok, why = plonk_verifier.verify(oracle, sc.circuit.as_dict(), pw_bad, aux_bad)
        assert not ok and "vanishing" in why


@pytest.mark.gpu
def test_device_proof_equals_the_oracle_proof(pkg, ctx, oracle):
    """Same witness, same transcript: the device proof is the oracle's proof, byte for byte."""
    public_inputs, sc = _pis_and_circuit(pkg, oracle, 7, 5)
    cfg = pkg.FriConfig(rate_bits=3, cap_height=2, proof_of_work_bits=6, num_query_rounds=5)
    pw_o, aux_o = _oracle_prove(pkg, oracle, sc, cfg, public_inputs)
    pw_d, aux_d = pkg.prove(ctx, sc, cfg, public_inputs)
    assert (aux_o["circuit_digest"] == aux_d["circuit_digest"]).all() and aux_o["widths"] == aux_d["widths"]
    assert pw_d.to_bytes(aux_d["fri_params"], aux_d["widths"]) == pw_o.to_bytes(aux_o["fri_params"], aux_o["widths"])
