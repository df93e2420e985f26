#!/usr/bin/env python3
"""Writes tests/golden/golden_vectors.json from the CPU oracle.

What these vectors are: FROZEN outputs of the in-repo restatement of the plonky2 algorithm (oracle/), so that
(a) the oracle cannot drift silently, (b) the CUDA path is checked against committed numbers and not only against
whatever the oracle computes today, and (c) a maintainer with a plonky2 checkout can diff them against the real
crate (the reference mount holds no source: SURVEY.md section 0).  Only `poseidon.zeros` is independently pinned
(the published plonky2 known-answer vector); everything else is the oracle's own output -- "parity unpinned".

Inputs are SplitMix64 streams reduced mod p (tests/conftest.py `splitmix_field`), so any implementation can
regenerate them from the seeds stored with each case.  Run:  python tests/golden/make_golden.py"""
import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from conftest import P, splitmix_field  # noqa: E402
from oracle import oracle as O  # noqa: E402


def hx(a):
    return ["%016x" % int(x) for x in np.asarray(a, dtype=np.uint64).reshape(-1)]


def digest(a):
    return hashlib.sha256(np.ascontiguousarray(a, dtype="<u8").tobytes()).hexdigest()


def arity_bits(degree_bits, rate_bits, cap_height, arity=4, final=5):
    out = []
    while degree_bits > final and degree_bits + rate_bits - arity >= cap_height:
        out.append(arity)
        degree_bits -= arity
    return out


def main():
    g = {"p": "%016x" % P, "note": "see make_golden.py; only poseidon.zeros is pinned by a published vector"}
    g["field"] = {"root_of_unity_32": "%016x" % O.lib().orc_root_of_unity(32),
                  "root_of_unity_20": "%016x" % O.lib().orc_root_of_unity(20),
                  "inv_7": "%016x" % O.lib().orc_inv(7), "mul_edge": "%016x" % O.lib().orc_mul(P - 1, P - 2)}
    g["poseidon"] = {"zeros": hx(O.poseidon(np.zeros(12, np.uint64))),
                     "iota": hx(O.poseidon(np.arange(12, dtype=np.uint64))),
                     "neg_one": hx(O.poseidon(np.full(12, P - 1, dtype=np.uint64))),
                     "seed_41": hx(O.poseidon(splitmix_field(41, (12,))))}
    x = splitmix_field(42, (8,))
    g["ntt"] = {"seed": 42, "n": 8, "fft": hx(O.fft(x)), "ifft": hx(O.ifft(x)), "coset_fft_7": hx(O.coset_fft(x, 7)),
                "lde_rate1": hx(O.lde(x, 1, 7))}
    x = splitmix_field(43, (1 << 12,))
    g["ntt_4096"] = {"seed": 43, "fft_sha256": digest(O.fft(x)), "ifft_sha256": digest(O.ifft(x)),
                     "coset_fft_sha256": digest(O.coset_fft(x, 7))}
    data = splitmix_field(44, (23,))
    g["hash_no_pad"] = {"seed": 44, "by_len": {str(k): hx(O.hash_no_pad(data[:k])) for k in (0, 1, 4, 5, 8, 9, 16, 17, 23)}}
    g["hash_or_noop"] = {"seed": 44, "by_len": {str(k): hx(O.hash_or_noop(data[:k])) for k in (0, 3, 4, 5)}}
    g["two_to_one"] = {"seed": 44, "out": hx(O.two_to_one(data[:4], data[4:8]))}
    g["merkle"] = []
    for log_leaves, leaf_len, cap_height in ((0, 3, 0), (3, 2, 0), (4, 7, 2), (6, 135, 4), (5, 4, 5)):
        leaves = splitmix_field(45 + log_leaves, (1 << log_leaves, leaf_len))
        digests, cap = O.merkle_build(leaves, cap_height)
        case = {"seed": 45 + log_leaves, "log_leaves": log_leaves, "leaf_len": leaf_len, "cap_height": cap_height,
                "cap": hx(cap), "digests_plonky2_sha256": digest(O.merkle_digests_to_plonky2(digests, log_leaves, cap_height))}
        if log_leaves > cap_height:
            case["proof_leaf_1"] = hx(O.merkle_prove(digests, log_leaves, cap_height, 1))
        g["merkle"].append(case)
    g["commit"] = []
    for log_n, n_cols, rate_bits, cap_height, from_coeffs, salted in (
            (3, 2, 1, 0, False, False), (5, 7, 2, 1, False, False), (8, 135, 3, 4, False, False),
            (10, 9, 3, 4, True, False), (6, 5, 1, 2, False, True), (12, 20, 3, 4, False, False)):
        seed = 0x5EED0000 + log_n
        vals = splitmix_field(seed, (n_cols, 1 << log_n))
        salt = splitmix_field(seed + 1, (4, 1 << (log_n + rate_bits))) if salted else None
        b = O.batch_from_values(vals, rate_bits, cap_height, salt=salt, from_coeffs=from_coeffs)
        z = splitmix_field(seed + 2, (2,))
        g["commit"].append({
            "seed": seed, "log_n": log_n, "n_cols": n_cols, "rate_bits": rate_bits, "cap_height": cap_height,
            "from_coeffs": from_coeffs, "salt_seed": seed + 1 if salted else None, "cap": hx(b["cap"]),
            "coeffs_sha256": digest(b["coeffs"]), "leaves_sha256": digest(b["leaves"]),
            "digests_plonky2_sha256": digest(O.merkle_digests_to_plonky2(b["digests"], log_n + rate_bits, cap_height)),
            "leaf_1": hx(b["leaves"][1]), "eval_point": hx(z), "eval_poly_0": hx(O.eval_poly_ext(b, 0, z, rate_bits, cap_height))})
    ch = O.Challenger()
    xs = splitmix_field(46, (29,))
    tr = []
    for lo, hi, k in ((0, 3, 2), (3, 11, 1), (11, 12, 9), (12, 29, 3)):
        ch.observe(xs[lo:hi])
        tr.append({"observe": [lo, hi], "challenges": hx([ch.get_challenge() for _ in range(k)])})
    g["challenger"] = {"seed": 46, "transcript": tr, "ext_challenge": hx(ch.get_extension_challenge())}
    g["fri"] = []
    for degree_bits, widths, rate_bits, cap_height, pow_bits, nq in ((5, (2, 2), 1, 0, 4, 3), (10, (5, 3, 2), 3, 2, 8, 5),
                                                                      (13, (9, 4), 1, 4, 10, 28)):
        batches = [O.batch_from_values(splitmix_field(47 + degree_bits + i, (w, 1 << degree_bits)), rate_bits, cap_height)
                   for i, w in enumerate(widths)]
        zeta = splitmix_field(48 + degree_bits, (2,))
        gen = O.lib().orc_root_of_unity(degree_bits)
        gz = np.array([O.lib().orc_mul(int(zeta[0]), gen), O.lib().orc_mul(int(zeta[1]), gen)], dtype=np.uint64)
        points = np.stack([zeta, gz])
        polys = [[(o, j) for o, w in enumerate(widths) for j in range(w)], [(0, j) for j in range(min(2, widths[0]))]]
        ch = O.Challenger()
        for b in batches:
            ch.observe(b["cap"].reshape(-1))
        ab = arity_bits(degree_bits, rate_bits, cap_height)
        proof = O.fri_prove_openings(batches, points, polys, ch, degree_bits, rate_bits, cap_height, pow_bits, nq, ab)
        g["fri"].append({"degree_bits": degree_bits, "widths": list(widths), "seed_base": 47 + degree_bits,
                         "zeta_seed": 48 + degree_bits, "rate_bits": rate_bits, "cap_height": cap_height,
                         "proof_of_work_bits": pow_bits, "num_query_rounds": nq, "arity_bits": ab,
                         "proof_words": int(proof.size), "proof_sha256": digest(proof), "first_cap_word": hx(proof[:1]),
                         "pow_witness": hx(proof[-1:]), "challenger_after": hx(list(ch.state))})
    # the prover steps either side of quotient evaluation (plonk/prover.rs): coset_ifft + chunks + commit, Zs + partial products
    x = splitmix_field(42, (8,))
    g["ntt"]["coset_ifft_7"] = hx(O.coset_ifft(x, 7))
    qv = splitmix_field(60, (2, 1 << 9))
    chunks = O.quotient_chunks(qv, 6, 8)
    qb = O.batch_from_values(chunks, 3, 2, from_coeffs=True)
    g["quotient"] = {"seed": 60, "degree_log": 6, "quotient_degree_bits": 3, "quotient_degree_factor": 8, "n_polys": 2,
                     "rate_bits": 3, "cap_height": 2, "chunks_sha256": digest(chunks), "chunk_0_head": hx(chunks[0][:4]),
                     "cap": hx(qb["cap"])}
    zw, zs_, zk = splitmix_field(61, (10, 64)), splitmix_field(62, (10, 64)), splitmix_field(63, (10,))
    zb, zg = splitmix_field(64, (2,)), splitmix_field(65, (2,))
    zz = O.partial_products_and_zs(zw, zs_, zk, zb, zg, 4)
    g["zs"] = {"wires_seed": 61, "sigmas_seed": 62, "k_is_seed": 63, "betas_seed": 64, "gammas_seed": 65, "log_n": 6,
               "n_routed": 10, "max_degree": 4, "shape": list(zz.shape), "sha256": digest(zz), "row_1": hx(zz[:, 1])}
    # a whole proof of a small synthetic circuit (16 rows, 20 wires of which 12 routed, NoopGate / ConstantGate(2) /
    # PublicInputGate / ArithmeticGate(3) in one selector group): the circuit data and the witness are stored in full, so
    # that a maintainer can replay them through plonky2's prove(); quotient values, caps and proof bytes are the oracle's
    import __graft_entry__ as GE
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import test_quotient as TQ
    pkg = GE.load_package()
    pis = splitmix_field(70, (5,))
    kw = dict(num_wires=20, num_routed_wires=12, num_ops=3, quotient_degree_factor=8)
    sc = pkg.synthetic_circuit(4, seed=71, public_inputs_hash=O.hash_no_pad(pis), **kw)
    cfg = pkg.FriConfig(rate_bits=3, cap_height=1, proof_of_work_bits=5, num_query_rounds=4)
    pw, aux = TQ._oracle_prove(pkg, O, sc, cfg, pis)
    c = sc.circuit
    g["plonk"] = {"degree_bits": 4, "config": kw, "rate_bits": 3, "cap_height": 1, "proof_of_work_bits": 5, "num_query_rounds": 4,
                  "gates": [[gt.kind, gt.param, gt.row, gt.selector_index, gt.group_start, gt.group_end] for gt in c.gates],
                  "public_inputs": hx(pis), "constants": hx(sc.constants), "sigmas": hx(sc.sigmas), "wires": hx(sc.wires),
                  "circuit_digest": hx(aux["circuit_digest"]), "wires_cap": hx(pw.proof.wires_cap),
                  "zs_cap": hx(pw.proof.plonk_zs_partial_products_cap), "quotient_cap": hx(pw.proof.quotient_polys_cap),
                  "openings_sha256": digest(np.concatenate([f.reshape(-1) for f in pw.proof.openings.fields()])),
                  "proof_bytes_sha256": hashlib.sha256(pw.to_bytes(aux["fri_params"], aux["widths"])).hexdigest(),
                  "proof_bytes_len": len(pw.to_bytes(aux["fri_params"], aux["widths"]))}
    # gates beyond the basic four: two PoseidonGate rows (gates/poseidon.rs wire map; row 1 with swap = 1) and the quotient
    # values of a 64-row circuit with every gate kind in four selector groups (synthetic_circuit(gate_set="all"))
    rows = np.zeros((2, 135), dtype=np.uint64)
    rows[:, :12] = splitmix_field(80, (2, 12))
    rows[1, 24] = 1
    O.poseidon_gate_fill(rows)
    g["poseidon_gate"] = {"inputs_seed": 80, "rows_sha256": digest(rows), "row0_outputs": hx(rows[0, 12:24]),
                          "row1_deltas": hx(rows[1, 25:29]), "row1_partial_sbox_inputs_head": hx(rows[1, 65:68]),
                          "row1_last_full_sbox_inputs_head": hx(rows[1, 123:126])}
    sa = pkg.synthetic_circuit(6, seed=81, gate_set="all", poseidon_fill=O.poseidon_gate_fill)
    be, ga, al = TQ._challenges(82, 2)
    ob = TQ._oracle_commitments(O, sa, be, ga)
    qa = O.quotient_values(sa.circuit.as_dict(), ob[0], ob[1], ob[2], sa.public_inputs_hash, be, ga, al)
    g["gates"] = {"degree_bits": 6, "seed": 81, "challenge_seed": 82,
                  "gates": [[gt.kind, gt.param, gt.row, gt.selector_index, gt.group_start, gt.group_end] for gt in sa.circuit.gates],
                  "wires_sha256": digest(sa.wires), "constants_sha256": digest(sa.constants), "sigmas_sha256": digest(sa.sigmas),
                  "public_inputs_hash": hx(sa.public_inputs_hash), "quotient_values_sha256": digest(qa),
                  "quotient_values_head": hx(qa[0, :4])}
    with open(os.path.join(HERE, "golden_vectors.json"), "w") as f:
        json.dump(g, f, indent=1)
    print("wrote golden_vectors.json: %d commit cases, %d fri cases" % (len(g["commit"]), len(g["fri"])))


if __name__ == "__main__":
    main()
