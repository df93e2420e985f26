"""Committed golden vectors (tests/golden/golden_vectors.json, written by tests/golden/make_golden.py).
CPU half: the oracle still reproduces them (and the one published plonky2 vector).  GPU half: the CUDA path
reproduces the same committed numbers through the C ABI -- not just whatever the oracle computes today."""
import hashlib
import json
import os

import numpy as np
import pytest

from conftest import P, ROOT, splitmix_field

G = json.load(open(os.path.join(ROOT, "tests", "golden", "golden_vectors.json")))
PUBLISHED_ZEROS = ("3c18a9786cb0b359 c4055e3364a246c3 7953db0ab48808f4 c71603f33a1144ca d7709673896996dc 46a84e87642f44ed "
                   "d032648251ee0b3c 1c687363b207df62 df8565563e8045fe 40f5b37ff4254dae d070f637b431067c 1792b1c4342109d7")


def hx(a):
    return ["%016x" % int(x) for x in np.asarray(a, dtype=np.uint64).reshape(-1)]


def digest(a):
    return hashlib.sha256(np.ascontiguousarray(a, dtype="<u8").tobytes()).hexdigest()


def _fri_inputs(c, make_batch, root_mul):
    batches = [make_batch(splitmix_field(c["seed_base"] + i, (w, 1 << c["degree_bits"]))) for i, w in enumerate(c["widths"])]
    zeta = splitmix_field(c["zeta_seed"], (2,))
    gz = np.array([root_mul(int(zeta[0])), root_mul(int(zeta[1]))], dtype=np.uint64)
    polys = [[(o, j) for o, w in enumerate(c["widths"]) for j in range(w)], [(0, j) for j in range(min(2, c["widths"][0]))]]
    return batches, np.stack([zeta, gz]), polys


# ------------------------------------------------------------------ CPU: oracle vs the committed vectors
def test_published_vector_is_in_the_fixture():
    assert G["poseidon"]["zeros"] == PUBLISHED_ZEROS.split()


def test_oracle_reproduces_golden_primitives(oracle):
    O = oracle
    assert hx(O.poseidon(np.zeros(12, np.uint64))) == G["poseidon"]["zeros"]
    assert hx(O.poseidon(np.arange(12, dtype=np.uint64))) == G["poseidon"]["iota"]
    assert hx(O.poseidon(np.full(12, P - 1, dtype=np.uint64))) == G["poseidon"]["neg_one"]
    assert hx(O.poseidon(splitmix_field(41, (12,)))) == G["poseidon"]["seed_41"]
    x = splitmix_field(G["ntt"]["seed"], (8,))
    assert hx(O.fft(x)) == G["ntt"]["fft"] and hx(O.ifft(x)) == G["ntt"]["ifft"]
    assert hx(O.coset_fft(x, 7)) == G["ntt"]["coset_fft_7"] and hx(O.lde(x, 1, 7)) == G["ntt"]["lde_rate1"]
    data = splitmix_field(44, (23,))
    for k, v in G["hash_no_pad"]["by_len"].items():
        assert hx(O.hash_no_pad(data[:int(k)])) == v
    for k, v in G["hash_or_noop"]["by_len"].items():
        assert hx(O.hash_or_noop(data[:int(k)])) == v
    assert hx(O.two_to_one(data[:4], data[4:8])) == G["two_to_one"]["out"]


def test_oracle_reproduces_golden_commits_and_fri(oracle):
    O = oracle
    for c in G["commit"]:
        vals = splitmix_field(c["seed"], (c["n_cols"], 1 << c["log_n"]))
        salt = splitmix_field(c["salt_seed"], (4, 1 << (c["log_n"] + c["rate_bits"]))) if c["salt_seed"] else None
        b = O.batch_from_values(vals, c["rate_bits"], c["cap_height"], salt=salt, from_coeffs=c["from_coeffs"])
        assert hx(b["cap"]) == c["cap"] and digest(b["leaves"]) == c["leaves_sha256"]
    c = G["fri"][1]
    gen = O.lib().orc_root_of_unity(c["degree_bits"])
    batches, points, polys = _fri_inputs(c, lambda v: O.batch_from_values(v, c["rate_bits"], c["cap_height"]),
                                         lambda x: O.lib().orc_mul(x, gen))
    ch = O.Challenger()
    for b in batches:
        ch.observe(b["cap"].reshape(-1))
    proof = O.fri_prove_openings(batches, points, polys, ch, c["degree_bits"], c["rate_bits"], c["cap_height"],
                                 c["proof_of_work_bits"], c["num_query_rounds"], c["arity_bits"])
    assert digest(proof) == c["proof_sha256"]


def _zs_inputs():
    z = G["zs"]
    n = 1 << z["log_n"]
    return (splitmix_field(z["wires_seed"], (z["n_routed"], n)), splitmix_field(z["sigmas_seed"], (z["n_routed"], n)),
            splitmix_field(z["k_is_seed"], (z["n_routed"],)), splitmix_field(z["betas_seed"], (2,)),
            splitmix_field(z["gammas_seed"], (2,)), z["max_degree"])


def test_oracle_reproduces_golden_quotient_and_zs(oracle):
    O = oracle
    assert hx(O.coset_ifft(splitmix_field(G["ntt"]["seed"], (8,)), 7)) == G["ntt"]["coset_ifft_7"]
    q = G["quotient"]
    chunks = O.quotient_chunks(splitmix_field(q["seed"], (q["n_polys"], 1 << (q["degree_log"] + q["quotient_degree_bits"]))),
                               q["degree_log"], q["quotient_degree_factor"])
    assert digest(chunks) == q["chunks_sha256"] and hx(chunks[0][:4]) == q["chunk_0_head"]
    assert hx(O.batch_from_values(chunks, q["rate_bits"], q["cap_height"], from_coeffs=True)["cap"]) == q["cap"]
    zz = O.partial_products_and_zs(*_zs_inputs())
    assert list(zz.shape) == G["zs"]["shape"] and digest(zz) == G["zs"]["sha256"] and hx(zz[:, 1]) == G["zs"]["row_1"]


def _golden_circuit(pkg):
    import __graft_entry__ as GE
    q = G["plonk"]
    un = lambda h, shape: np.array([int(x, 16) for x in h], dtype=np.uint64).reshape(shape)
    n = 1 << q["degree_bits"]
    kw = q["config"]
    gates = [pkg.Gate(*g) for g in q["gates"]]
    circ = pkg.CircuitDescription(q["degree_bits"], gates, rate_bits=q["rate_bits"], quotient_degree_factor=kw["quotient_degree_factor"],
                                  num_wires=kw["num_wires"], num_routed_wires=kw["num_routed_wires"])
    pis = un(q["public_inputs"], (-1,))
    sc = pkg.SyntheticCircuit(circ, un(q["constants"], (3, n)), un(q["sigmas"], (kw["num_routed_wires"], n)),
                              un(q["wires"], (kw["num_wires"], n)), None)
    cfg = pkg.FriConfig(rate_bits=q["rate_bits"], cap_height=q["cap_height"], proof_of_work_bits=q["proof_of_work_bits"],
                        num_query_rounds=q["num_query_rounds"])
    return sc, cfg, pis


def _check_golden_proof(pw, aux):
    import hashlib
    q = G["plonk"]
    assert hx(aux["circuit_digest"]) == q["circuit_digest"]
    assert hx(pw.proof.wires_cap) == q["wires_cap"] and hx(pw.proof.plonk_zs_partial_products_cap) == q["zs_cap"]
    assert hx(pw.proof.quotient_polys_cap) == q["quotient_cap"]
    assert digest(np.concatenate([f.reshape(-1) for f in pw.proof.openings.fields()])) == q["openings_sha256"]
    data = pw.to_bytes(aux["fri_params"], aux["widths"])
    assert len(data) == q["proof_bytes_len"] and hashlib.sha256(data).hexdigest() == q["proof_bytes_sha256"]


def test_oracle_reproduces_golden_proof(oracle):
    """The whole prove() flow of a stored circuit + witness on the oracle: caps, opening set and ProofWithPublicInputs bytes."""
    import __graft_entry__ as GE
    import plonk_verifier
    import test_quotient as TQ
    pkg = GE.load_package()
    sc, cfg, pis = _golden_circuit(pkg)
    sc.public_inputs_hash = oracle.hash_no_pad(pis)
    pw, aux = TQ._oracle_prove(pkg, oracle, sc, cfg, pis)
    _check_golden_proof(pw, aux)
    assert plonk_verifier.verify(oracle, sc.circuit.as_dict(), pw, aux)[0]


def _golden_gate_rows():
    rows = np.zeros((2, 135), dtype=np.uint64)
    rows[:, :12] = splitmix_field(G["poseidon_gate"]["inputs_seed"], (2, 12))
    rows[1, 24] = 1
    return rows


def _check_golden_gate_rows(rows):
    q = G["poseidon_gate"]
    assert digest(rows) == q["rows_sha256"] and hx(rows[0, 12:24]) == q["row0_outputs"] and hx(rows[1, 25:29]) == q["row1_deltas"]
    assert hx(rows[1, 65:68]) == q["row1_partial_sbox_inputs_head"] and hx(rows[1, 123:126]) == q["row1_last_full_sbox_inputs_head"]


def _golden_gates_circuit(pkg, fill):
    q = G["gates"]
    sc = pkg.synthetic_circuit(q["degree_bits"], seed=q["seed"], gate_set="all", poseidon_fill=fill)
    assert [[g.kind, g.param, g.row, g.selector_index, g.group_start, g.group_end] for g in sc.circuit.gates] == q["gates"]
    assert digest(sc.wires) == q["wires_sha256"] and digest(sc.constants) == q["constants_sha256"]
    assert digest(sc.sigmas) == q["sigmas_sha256"] and hx(sc.public_inputs_hash) == q["public_inputs_hash"]
    return sc


def test_oracle_reproduces_golden_gates(oracle):
    """PoseidonGate rows and the quotient values of the circuit with every gate kind."""
    import __graft_entry__ as GE
    import test_quotient as TQ
    _check_golden_gate_rows(oracle.poseidon_gate_fill(_golden_gate_rows()))
    sc = _golden_gates_circuit(GE.load_package(), oracle.poseidon_gate_fill)
    be, ga, al = TQ._challenges(G["gates"]["challenge_seed"], 2)
    ob = TQ._oracle_commitments(oracle, sc, be, ga)
    qa = oracle.quotient_values(sc.circuit.as_dict(), ob[0], ob[1], ob[2], sc.public_inputs_hash, be, ga, al)
    assert digest(qa) == G["gates"]["quotient_values_sha256"] and hx(qa[0, :4]) == G["gates"]["quotient_values_head"]


# ------------------------------------------------------------------ GPU: the CUDA path vs the committed vectors
@pytest.fixture(scope="module")
def pkg():
    import __graft_entry__ as GE
    return GE.load_package()


@pytest.fixture(scope="module")
def ctx(pkg):
    c = pkg.Context(0)
    yield c
    c.close()


@pytest.mark.gpu
def test_device_reproduces_golden_primitives(pkg, ctx):
    st = np.stack([np.zeros(12, np.uint64), np.arange(12, dtype=np.uint64), np.full(12, P - 1, dtype=np.uint64),
                   splitmix_field(41, (12,))])
    out = ctx.poseidon_permute(st)
    for row, key in zip(out, ("zeros", "iota", "neg_one", "seed_41")):
        assert hx(row) == G["poseidon"][key]
    x = splitmix_field(G["ntt"]["seed"], (1, 8))
    assert hx(ctx.ntt(x, "fft")) == G["ntt"]["fft"] and hx(ctx.ntt(x, "ifft")) == G["ntt"]["ifft"]
    assert hx(ctx.ntt(x, "coset_fft", 7)) == G["ntt"]["coset_fft_7"]
    x = splitmix_field(G["ntt_4096"]["seed"], (1, 1 << 12))
    assert digest(ctx.ntt(x, "fft")) == G["ntt_4096"]["fft_sha256"]
    assert digest(ctx.ntt(x, "ifft")) == G["ntt_4096"]["ifft_sha256"]
    assert digest(ctx.ntt(x, "coset_fft", 7)) == G["ntt_4096"]["coset_fft_sha256"]
    for c in G["merkle"]:
        leaves = splitmix_field(c["seed"], (1 << c["log_leaves"], c["leaf_len"]))
        t = pkg.MerkleTree(ctx, leaves, c["cap_height"])
        assert hx(t.cap) == c["cap"]
        assert digest(t.digests) == c["digests_plonky2_sha256"]
        if "proof_leaf_1" in c:
            assert hx(t.prove(1)) == c["proof_leaf_1"]
        t.close()


@pytest.mark.gpu
def test_device_reproduces_golden_quotient_and_zs(pkg, ctx):
    assert hx(ctx.ntt(splitmix_field(G["ntt"]["seed"], (1, 8)), "coset_ifft", 7)) == G["ntt"]["coset_ifft_7"]
    q = G["quotient"]
    vals = splitmix_field(q["seed"], (q["n_polys"], 1 << (q["degree_log"] + q["quotient_degree_bits"])))
    b = pkg.PolynomialBatch.from_quotient_values(ctx, vals, q["degree_log"], q["quotient_degree_factor"], q["rate_bits"],
                                                 False, q["cap_height"])
    assert hx(b.cap) == q["cap"] and digest(b.polynomials) == q["chunks_sha256"]
    b.close()
    zz = ctx.partial_products_and_zs(*_zs_inputs())
    assert list(zz.shape) == G["zs"]["shape"] and digest(zz) == G["zs"]["sha256"] and hx(zz[:, 1]) == G["zs"]["row_1"]


@pytest.mark.gpu
def test_device_reproduces_golden_commits(pkg, ctx):
    for c in G["commit"]:
        vals = splitmix_field(c["seed"], (c["n_cols"], 1 << c["log_n"]))
        salt = splitmix_field(c["salt_seed"], (4, 1 << (c["log_n"] + c["rate_bits"]))) if c["salt_seed"] else None
        mk = pkg.PolynomialBatch.from_coeffs if c["from_coeffs"] else pkg.PolynomialBatch.from_values
        b = mk(ctx, vals, c["rate_bits"], salt is not None, c["cap_height"], salt=salt)
        ex = b.export()
        assert hx(b.cap) == c["cap"]
        assert digest(ex["coeffs"]) == c["coeffs_sha256"] and digest(ex["leaves"]) == c["leaves_sha256"]
        assert digest(ex["digests"]) == c["digests_plonky2_sha256"]
        assert hx(ex["leaves"][1]) == c["leaf_1"]
        z = np.array([int(v, 16) for v in c["eval_point"]], dtype=np.uint64)
        assert hx(b.eval_commitment(z)[0]) == c["eval_poly_0"]
        b.close()


@pytest.mark.gpu
def test_device_reproduces_golden_challenger_and_fri(pkg, ctx):
    ch = pkg.Challenger(ctx)
    xs = splitmix_field(G["challenger"]["seed"], (29,))
    for step in G["challenger"]["transcript"]:
        lo, hi = step["observe"]
        ch.observe_elements(xs[lo:hi])
        assert hx(ch.get_n_challenges(len(step["challenges"]))) == step["challenges"]
    assert hx(ch.get_extension_challenge()) == G["challenger"]["ext_challenge"]
    for c in G["fri"]:
        w = pow(7, (P - 1) >> c["degree_bits"], P)
        batches, points, polys = _fri_inputs(
            c, lambda v: pkg.PolynomialBatch.from_values(ctx, v, c["rate_bits"], False, c["cap_height"]),
            lambda x: x * w % P)
        ch = pkg.Challenger(ctx)
        for b in batches:
            ch.observe_cap(b.cap)
        cfg = pkg.FriConfig(rate_bits=c["rate_bits"], cap_height=c["cap_height"], proof_of_work_bits=c["proof_of_work_bits"],
                            num_query_rounds=c["num_query_rounds"])
        fp = cfg.fri_params(c["degree_bits"])
        assert list(fp.reduction_arity_bits) == c["arity_bits"]
        proof = pkg.prove_openings(pkg.FriInstanceInfo(points, polys), batches, ch, fp)
        assert proof.size == c["proof_words"] and digest(proof) == c["proof_sha256"]
        assert hx(proof[-1:]) == c["pow_witness"]
        assert hx(ch.snapshot()[0]) == c["challenger_after"]
        for b in batches:
            b.close()


@pytest.mark.gpu
def test_device_reproduces_golden_proof(pkg, ctx):
    sc, cfg, pis = _golden_circuit(pkg)
    sc.public_inputs_hash = pkg.hash_no_pad(ctx, pis)
    pw, aux = pkg.prove(ctx, sc, cfg, pis)
    _check_golden_proof(pw, aux)


@pytest.mark.gpu
def test_device_reproduces_golden_gates(pkg, ctx):
    import test_quotient as TQ
    _check_golden_gate_rows(ctx.poseidon_gate_witness(_golden_gate_rows()))
    sc = _golden_gates_circuit(pkg, ctx.poseidon_gate_witness)
    be, ga, al = TQ._challenges(G["gates"]["challenge_seed"], 2)
    gb = TQ._device_commitments(pkg, ctx, sc, be, ga)
    qa = pkg.quotient_values(ctx, sc.circuit, gb[0], gb[1], gb[2], sc.public_inputs_hash, be, ga, al)
    assert digest(qa) == G["gates"]["quotient_values_sha256"] and hx(qa[0, :4]) == G["gates"]["quotient_values_head"]
    for b in gb:
        b.close()
