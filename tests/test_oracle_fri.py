"""Oracle tier for the FRI half of the path: challenger semantics and prover/verifier consistency."""
import numpy as np
import pytest

from conftest import P, splitmix_field


def test_challenger_semantics(oracle):
    ch = oracle.Challenger()
    # get_challenge with empty buffers duplexes the zero state: outputs pop from the BACK of state[0..8]
    z = oracle.poseidon(np.zeros(12, dtype=np.uint64))
    assert ch.get_challenge() == int(z[7]) and ch.get_challenge() == int(z[6])
    # observing invalidates buffered outputs; 8 inputs trigger a duplexing immediately
    ch = oracle.Challenger()
    xs = splitmix_field(3, (8,))
    ch.observe(xs[:3])
    assert ch.snapshot()[1] == [int(v) for v in xs[:3]] and ch.snapshot()[2] == []
    ch.observe(xs[3:])
    st = oracle.poseidon(np.concatenate([xs, np.zeros(4, dtype=np.uint64)]))
    assert ch.snapshot()[1] == [] and ch.snapshot()[2] == [int(v) for v in st[:8]]
    assert ch.get_challenge() == int(st[7])
    # overwrite mode: a short input overwrites only its own lanes
    ch.observe(xs[:2])
    e = ch.get_extension_challenge()
    st2 = st.copy()
    st2[:2] = xs[:2]
    st2 = oracle.poseidon(st2)
    assert [int(e[0]), int(e[1])] == [int(st2[7]), int(st2[6])]


def _setup(oracle, degree_bits, rate_bits, cap_height, widths, seed=5):
    batches = []
    for i, w in enumerate(widths):
        vals = splitmix_field(seed + i, (w, 1 << degree_bits))
        batches.append(oracle.batch_from_values(vals, rate_bits, cap_height))
    # two opening points in the extension field, like zeta and g * zeta
    zeta = splitmix_field(seed + 100, (2,))
    g = oracle.lib().orc_root_of_unity(degree_bits)
    gz = np.array([oracle.lib().orc_mul(int(zeta[0]), g), oracle.lib().orc_mul(int(zeta[1]), g)], dtype=np.uint64)
    all_polys = [(o, j) for o, w in enumerate(widths) for j in range(w)]
    polys = [all_polys, [(len(widths) - 1, j) for j in range(min(2, widths[-1]))]]
    points = np.stack([zeta, gz])
    openings = np.array([oracle.eval_poly_ext(batches[o], j, points[b]) for b, pl in enumerate(polys)
                         for (o, j) in pl], dtype=np.uint64)
    return batches, points, polys, openings


@pytest.mark.parametrize("degree_bits,rate_bits,cap_height,arity", [(6, 3, 1, [4]), (9, 3, 4, [4]), (10, 1, 2, [4, 4]),
                                                                   (5, 2, 0, [])])
def test_fri_prover_accepted_by_verifier(oracle, degree_bits, rate_bits, cap_height, arity):
    batches, points, polys, openings = _setup(oracle, degree_bits, rate_bits, cap_height, [3, 2])
    ch = oracle.Challenger()
    for b in batches:
        ch.observe(b["cap"])
    ch_v = ch.copy()
    proof = oracle.fri_prove_openings(batches, points, polys, ch, degree_bits, rate_bits, cap_height, 4, 5, arity)
    args = (degree_bits, rate_bits, cap_height, 4, 5, arity)
    assert oracle.fri_verify(batches, points, polys, openings, ch_v.copy(), *args, proof)
    # the transcript ends in the same state on both sides
    v = ch_v.copy()
    oracle.fri_verify(batches, points, polys, openings, v, *args, proof)
    assert v.snapshot() == ch.snapshot()
    # any single-word corruption is rejected; so is a wrong opening
    rng = np.random.default_rng(0)
    for pos in rng.choice(proof.size, size=12, replace=False):
        bad = proof.copy()
        bad[pos] = (int(bad[pos]) + 1) % P
        assert not oracle.fri_verify(batches, points, polys, openings, ch_v.copy(), *args, bad), pos
    bad_open = openings.copy()
    bad_open[0, 0] = (int(bad_open[0, 0]) + 1) % P
    assert not oracle.fri_verify(batches, points, polys, bad_open, ch_v.copy(), *args, proof)


def test_final_poly_is_low_degree_quotient(oracle):
    """final_poly = X * sum_b alpha^(..) (F_b(X) - F_b(z_b)) / (X - z_b): check by evaluation."""
    degree_bits = 5
    batches, points, polys, openings = _setup(oracle, degree_bits, 1, 0, [2, 2])
    alpha = splitmix_field(9, (2,))
    fin = oracle.fri_final_poly(batches, 1, 0, points, polys, alpha)
    assert fin.shape == (32, 2) and (fin[0] == 0).all()
    L = oracle.lib()

    def emul(a, b):
        return ((a[0] * b[0] + 7 * a[1] * b[1]) % P, (a[0] * b[1] + a[1] * b[0]) % P)

    def einv(a):
        n = pow((a[0] * a[0] - 7 * a[1] * a[1]) % P, P - 2, P)
        return (a[0] * n % P, (-a[1]) * n % P)

    x = (123456789, 987654321)
    # evaluate final poly at x
    acc = (0, 0)
    for c in reversed(fin):
        acc = emul(acc, x)
        acc = ((acc[0] + int(c[0])) % P, (acc[1] + int(c[1])) % P)
    al = (int(alpha[0]), int(alpha[1]))
    total, pos = (0, 0), 0
    for b, pl in enumerate(polys):
        red_x, red_z, ap = (0, 0), (0, 0), (1, 0)
        for (o, j) in pl:
            fx = oracle.eval_poly_ext(batches[o], j, np.array(x, dtype=np.uint64))
            fz = openings[pos]
            red_x = tuple((red_x[k] + emul(ap, (int(fx[0]), int(fx[1])))[k]) % P for k in range(2))
            red_z = tuple((red_z[k] + emul(ap, (int(fz[0]), int(fz[1])))[k]) % P for k in range(2))
            ap = emul(ap, al)
            pos += 1
        z = (int(points[b][0]), int(points[b][1]))
        q = emul(((red_x[0] - red_z[0]) % P, (red_x[1] - red_z[1]) % P), einv(((x[0] - z[0]) % P, (x[1] - z[1]) % P)))
        sh = (1, 0)
        for _ in pl:
            sh = emul(sh, al)
        total = emul(total, sh)
        total = ((total[0] + q[0]) % P, (total[1] + q[1]) % P)
    assert acc == emul(total, x)


def test_fri_proof_struct_and_bytes_roundtrip(oracle):
    """Host-side wire format of a FriProof (serialization.py) on an oracle proof: words -> struct -> bytes -> words,
    and the struct's Merkle openings verify against the caps."""
    import __graft_entry__ as G
    pkg = G.load_package()
    ser = __import__(pkg.__name__ + ".serialization", fromlist=["x"])
    degree_bits, rate_bits, cap_height, pow_bits, nq, widths = 10, 2, 1, 5, 4, (3, 2)
    batches = [oracle.batch_from_values(splitmix_field(90 + i, (w, 1 << degree_bits)), rate_bits, cap_height)
               for i, w in enumerate(widths)]
    points = splitmix_field(95, (1, 2))
    polys = [[(o, j) for o, w in enumerate(widths) for j in range(w)]]
    ch = oracle.Challenger()
    arity = [4, 4]
    words = oracle.fri_prove_openings(batches, points, polys, ch, degree_bits, rate_bits, cap_height, pow_bits, nq, arity)
    st = ser.fri_proof_struct(words, widths, degree_bits, rate_bits, cap_height, arity, nq)
    assert len(st["commit_phase_merkle_caps"]) == 2 and len(st["query_round_proofs"]) == nq
    assert st["final_poly"].shape == (1 << (degree_bits - 8), 2) and st["pow_witness"] == int(words[-1])
    log_nb = degree_bits + rate_bits
    for qr in st["query_round_proofs"]:
        for (leaf, sib), b in zip(qr["initial_trees_proof"], batches):
            # the opened leaf is one of the committed rows and its path verifies against the oracle's cap
            idx = np.where((b["leaves"] == leaf).all(axis=1))[0]
            assert idx.size >= 1 and oracle.merkle_verify(leaf, int(idx[0]), sib, b["cap"])
        assert [e.shape for e, _ in qr["steps"]] == [(16, 2), (16, 2)]
        assert [s.shape[0] for _, s in qr["steps"]] == [log_nb - 4 - cap_height, log_nb - 8 - cap_height]
    data = ser.write_fri_proof(st)
    n_paths = nq * (len(widths) + len(arity))
    assert len(data) == 8 * words.size + n_paths
    back = ser.read_fri_proof(data, widths, degree_bits, rate_bits, cap_height, arity, nq)
    assert (back == words).all()
    for bad in (data[:-1], data + b"\0", data[:40]):
        with pytest.raises(ValueError):
            ser.read_fri_proof(bad, widths, degree_bits, rate_bits, cap_height, arity, nq)
    with pytest.raises(ValueError):
        ser.fri_proof_struct(words[:-1], widths, degree_bits, rate_bits, cap_height, arity, nq)


def test_splice_initial_openings_rebuilds_the_proof(oracle):
    """serialization.splice_initial_openings (sharded prove): FRI-only stream + per-oracle openings -> the full stream."""
    import __graft_entry__ as G
    pkg = G.load_package()
    ser = __import__(pkg.__name__ + ".serialization", fromlist=["x"])
    degree_bits, rate_bits, cap_height, nq, widths, arity = 9, 1, 2, 3, (4, 2), [4]
    batches = [oracle.batch_from_values(splitmix_field(120 + i, (w, 1 << degree_bits)), rate_bits, cap_height)
               for i, w in enumerate(widths)]
    polys = [[(o, j) for o, w in enumerate(widths) for j in range(w)]]
    words = oracle.fri_prove_openings(batches, splitmix_field(125, (1, 2)), polys, oracle.Challenger(), degree_bits,
                                      rate_bits, cap_height, 3, nq, arity)
    st = ser.fri_proof_struct(words, widths, degree_bits, rate_bits, cap_height, arity, nq)
    fri_only = [np.concatenate([c.reshape(-1) for c in st["commit_phase_merkle_caps"]])]
    for qr in st["query_round_proofs"]:
        for evals, sib in qr["steps"]:
            fri_only += [evals.reshape(-1), sib.reshape(-1)]
    fri_only += [st["final_poly"].reshape(-1), np.array([st["pow_witness"]], dtype=np.uint64)]
    openings = [(np.stack([qr["initial_trees_proof"][o][0] for qr in st["query_round_proofs"]]),
                 np.stack([qr["initial_trees_proof"][o][1] for qr in st["query_round_proofs"]])) for o in range(len(widths))]
    back = ser.splice_initial_openings(np.concatenate(fri_only), openings, degree_bits, rate_bits, cap_height, arity, nq)
    assert (back == words).all()
    with pytest.raises(ValueError):
        ser.splice_initial_openings(np.concatenate(fri_only)[:-1], openings, degree_bits, rate_bits, cap_height, arity, nq)
