"""The drop-in boundary driven from plain C (tests/c_harness/abi_harness.c): every record the C program prints is
recomputed with the oracle -- commit, openings, Merkle paths, the opening set and a whole FRI proof."""
import subprocess

import numpy as np
import pytest

from conftest import P, splitmix_field
import c_harness_build

pytestmark = pytest.mark.gpu


def _words(tok):
    return np.array([int(t, 16) for t in tok], dtype=np.uint64)


@pytest.mark.parametrize("log_n,n_cols,rate_bits,cap_height,seed", [(10, 5, 3, 4, 11), (13, 20, 1, 0, 12), (6, 1, 2, 2, 13)])
def test_c_program_matches_oracle(oracle, log_n, n_cols, rate_bits, cap_height, seed):
    exe = c_harness_build.build()
    r = subprocess.run([exe, str(log_n), str(n_cols), str(rate_bits), str(cap_height), str(seed)], capture_output=True,
                       text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    rec = [ln.split() for ln in r.stdout.strip().splitlines()]
    vals = splitmix_field(seed, (n_cols, 1 << log_n))  # same SplitMix64 stream as the C program
    want = oracle.batch_from_values(vals, rate_bits, cap_height)
    log_nb = log_n + rate_bits
    by = {}
    for t in rec:
        by.setdefault(t[0], []).append(t[1:])
    assert (_words(by["cap"][0]).reshape(-1, 4) == want["cap"]).all()
    assert len(by["row"]) == 3 and len(by["path"]) == 3
    for row, path in zip(by["row"], by["path"]):
        leaf = int(row[0])
        assert int(path[0]) == leaf
        assert (_words(row[1:]) == want["leaves"][leaf]).all()
        sib = _words(path[1:]).reshape(-1, 4)
        if want["digests"] is not None and sib.size:
            assert (sib == np.asarray(oracle.merkle_prove(want["digests"], log_nb, cap_height, leaf)).reshape(-1, 4)).all()
        assert oracle.merkle_verify(want["leaves"][leaf], leaf, sib, want["cap"])
    z = np.array([seed % P, (seed + 1) % P], dtype=np.uint64)
    ev = _words(by["eval"][0]).reshape(n_cols, 2)
    for j in range(n_cols):
        assert (ev[j] == oracle.eval_poly_ext(want, j, z, rate_bits, cap_height)).all()
    # the FRI proof: same challenger transcript as the C program
    ch = oracle.Challenger()
    ch.observe(want["cap"].reshape(-1))
    ch.observe(ev.reshape(-1))
    arity, db = [], log_n
    while db > 5 and db + rate_bits - 4 >= cap_height:
        arity.append(4)
        db -= 4
    points, polys = z.reshape(1, 2), [[(0, j) for j in range(n_cols)]]
    proof = oracle.fri_prove_openings([want], points, polys, ch, log_n, rate_bits, cap_height, 8, 4, arity)
    assert (_words(by["proof"][0]) == proof).all()
    # Zs / partial products and the quotient hand-over, same SplitMix streams as the C program
    n = 1 << log_n
    n_routed = min(n_cols, 5)
    zs = oracle.partial_products_and_zs(splitmix_field(seed, (n_routed, n)), splitmix_field(seed + 100, (n_routed, n)),
                                        splitmix_field(seed + 200, (n_routed,)), [seed % P + 2, seed % P + 3],
                                        [seed % P + 4, seed % P + 5], 2)
    assert len(by["zs"]) == zs.shape[0]
    for t in by["zs"]:
        assert int(t[1], 16) == int(zs[int(t[0]), 1 % n])
    fnv = 0
    for v in zs.reshape(-1):
        fnv = ((fnv * 0x100000001B3) ^ int(v)) & 0xFFFFFFFFFFFFFFFF
    assert int(by["zs_fnv"][0][0], 16) == fnv
    q_log_n = min(log_n, 10)
    chunks = oracle.quotient_chunks(splitmix_field(seed + 300, (2, 1 << (q_log_n + 3))), q_log_n, 8)
    qcap = oracle.batch_from_values(chunks, rate_bits, 0, from_coeffs=True, want=())["cap"]
    assert int(by["qcap"][0][0]) == q_log_n and (_words(by["qcap"][0][1:]).reshape(-1, 4) == qcap).all()
    # the sharded commit and proof through bsx_comm (2 and 8 emulated ranks) equal the single-GPU ones word for word
    assert [t for t in by["comm"]] == [["2", "1", "1", "1"], ["8", "1", "1", "1"]]
    # bytes of the proof stream: the C ABI writer against the Python writer over the oracle's proof
    import __graft_entry__ as G
    pkg = G.load_package()
    ser = __import__(pkg.__name__ + ".serialization", fromlist=["x"])
    data = ser.write_fri_proof(ser.fri_proof_struct(proof, [n_cols], log_n, rate_bits, cap_height, arity, 4))
    fnv = 0xcbf29ce484222325
    for b in data:
        fnv = ((fnv ^ b) * 0x100000001B3) & 0xFFFFFFFFFFFFFFFF
    fb = by["fribytes"][0]
    assert int(fb[0]) == len(data) and int(fb[1], 16) == fnv and int(fb[2]) == proof.size and int(fb[3]) == int(proof[-1])
    # quotient evaluation over the fourteen-kind gate table and the PoseidonGate generator: same streams as the C program
    qn = min(log_n, 7)
    kinds = [(0, 0, 0, 0, 0, 6), (1, 2, 1, 0, 0, 6), (2, 0, 2, 0, 0, 6), (12, 0, 3, 0, 0, 6), (7, 32, 4, 0, 0, 6), (9, 43, 5, 0, 0, 6),
             (10, 32, 6, 1, 6, 11), (3, 20, 7, 1, 6, 11), (5, 10, 8, 1, 6, 11), (6, 13, 9, 1, 6, 11), (8, 66, 10, 1, 6, 11),
             (11, 4 | 4 << 8 | 2 << 16, 11, 2, 11, 13), (13, 4 | 6 << 8, 12, 2, 11, 13), (4, 0, 13, 3, 13, 14)]
    circ = {"degree_bits": qn, "rate_bits": 3, "quotient_degree_factor": 8, "num_challenges": 2, "num_wires": 135,
            "num_routed_wires": 80, "num_constants": 6, "num_selectors": 4, "num_partial_products": 9, "num_gate_constraints": 123,
            "gates": kinds, "k_is": splitmix_field(seed + 400, (80,))}
    bs = [oracle.batch_from_values(splitmix_field(seed + 500 + k, (w, 1 << qn)), 3, 0) for k, w in enumerate((86, 135, 20))]
    s0 = seed % P
    qv = oracle.quotient_values(circ, bs[0], bs[1], bs[2], [s0, 1, 2, 3], [s0 + 6, s0 + 7], [s0 + 8, s0 + 9], [s0 + 10, s0 + 11])
    fnv = 0
    for v in qv.reshape(-1):
        fnv = ((fnv * 0x100000001B3) ^ int(v)) & 0xFFFFFFFFFFFFFFFF
    t = by["qv"][0]
    assert int(t[0]) == qn and int(t[1], 16) == fnv and int(t[2], 16) == int(qv[0, 0]) and int(t[3], 16) == int(qv[-1, -1])
    rows = splitmix_field(seed + 600, (2, 135))
    rows[0, 24], rows[1, 24] = 0, 1
    assert (_words(by["pgw"][0]).reshape(2, 135) == oracle.poseidon_gate_fill(rows)).all()
    misuse = by["misuse"][0]
    assert int(misuse[0]) == -1 and "cap_height=5 should be at most log2(leaves.len())=1" in " ".join(misuse[1:])
