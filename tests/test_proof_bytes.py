"""`ProofWithPublicInputs::to_bytes` and the typed view of the FriProof stream: three
independently written writers -- the C ABI (csrc/api_proof.cu), the oracle's C writer (plonky2_plonk_oracle.c) and
the Python composition over serialization.write_fri_proof -- must agree byte for byte on a proof produced by the
oracle's prover; from_bytes restores every part.  Host code only (no GPU)."""
import numpy as np
import pytest

from conftest import splitmix_field
from test_oracle_fri import _setup
import __graft_entry__ as G


@pytest.fixture(scope="module")
def pkg():
    return G.load_package()


@pytest.mark.parametrize("degree_bits,rate_bits,cap_height,widths,n_pis", [
    (9, 3, 4, [5, 7, 4, 3], 4), (10, 1, 2, [3, 2, 2, 1], 0), (6, 2, 0, [2, 2, 2, 2], 9)])
def test_three_writers_agree(pkg, oracle, degree_bits, rate_bits, cap_height, widths, n_pis):
    ser = __import__(pkg.__name__ + ".serialization", fromlist=["x"])
    batches, points, polys, _ = _setup(oracle, degree_bits, rate_bits, cap_height, widths)
    cfg = pkg.FriConfig(rate_bits=rate_bits, cap_height=cap_height, proof_of_work_bits=5, num_query_rounds=6)
    fp = cfg.fri_params(degree_bits)
    arity = list(fp.reduction_arity_bits)
    ch = oracle.Challenger()
    words = oracle.fri_prove_openings(batches, points, polys, ch, degree_bits, rate_bits, cap_height, 5, 6, arity)
    # an opening set of the right shape (values are arbitrary here: the writers only move them)
    n_open = [2, widths[0] - 2, widths[1], 2, 2, widths[2] - 2, widths[3]]
    ops = [splitmix_field(70 + k, (max(n, 0), 2)) for k, n in enumerate(n_open)]
    pis = splitmix_field(99, (n_pis,))
    caps = [batches[1]["cap"], batches[2]["cap"], batches[3]["cap"]]
    pw = pkg.ProofWithPublicInputs(pkg.Proof(caps[0], caps[1], caps[2], pkg.OpeningSet(*ops), words), pis)
    got = pw.to_bytes(fp, widths)
    want_oracle = oracle.write_proof_with_public_inputs(caps[0], caps[1], caps[2], ops, words, degree_bits, rate_bits,
                                                        cap_height, 5, 6, arity, widths, pis)
    st = ser.fri_proof_struct(words, widths, degree_bits, rate_bits, cap_height, arity, 6)
    want_py = b"".join([np.asarray(c).astype("<u8").tobytes() for c in caps] +
                       [o.astype("<u8").tobytes() for o in ops] + [ser.write_fri_proof(st), pis.astype("<u8").tobytes()])
    assert got == want_oracle == want_py
    assert pkg.fri_proof_bytes(words, fp, widths) == ser.write_fri_proof(st)
    back = pkg.ProofWithPublicInputs.from_bytes(got, fp, widths, n_open)
    assert (back.proof.opening_proof == words).all() and (back.public_inputs == pis).all()
    assert (back.proof.quotient_polys_cap == caps[2]).all() and (back.proof.openings.wires == ops[2]).all()
    assert back.to_bytes(fp, widths) == got
    with pytest.raises(ValueError):
        pkg.ProofWithPublicInputs.from_bytes(got[:-3], fp, widths, n_open)


def test_layout_matches_struct_view(pkg, oracle):
    ser = __import__(pkg.__name__ + ".serialization", fromlist=["x"])
    degree_bits, rate_bits, cap_height, widths = 10, 2, 1, [4, 3]
    batches, points, polys, _ = _setup(oracle, degree_bits, rate_bits, cap_height, widths)
    fp = pkg.FriConfig(rate_bits=rate_bits, cap_height=cap_height, proof_of_work_bits=4, num_query_rounds=3).fri_params(degree_bits)
    arity = list(fp.reduction_arity_bits)
    words = oracle.fri_prove_openings(batches, points, polys, oracle.Challenger(), degree_bits, rate_bits, cap_height, 4, 3, arity)
    lay = pkg.fri_proof_layout(fp, widths)
    st = ser.fri_proof_struct(words, widths, degree_bits, rate_bits, cap_height, arity, 3)
    assert lay.total_words == words.size and lay.n_rounds == len(arity) and lay.n_oracles == 2
    for r in range(len(arity)):
        o = lay.caps_offset + r * lay.cap_words
        assert (words[o:o + lay.cap_words].reshape(-1, 4) == st["commit_phase_merkle_caps"][r]).all()
    for q in range(3):
        base = lay.query_offset + q * lay.query_stride
        for o in range(2):
            leaf, sib = st["query_round_proofs"][q]["initial_trees_proof"][o]
            a = base + lay.initial_leaf_offset[o]
            assert (words[a:a + lay.initial_leaf_words[o]] == leaf).all()
            a = base + lay.initial_proof_offset[o]
            assert (words[a:a + 4 * lay.initial_siblings].reshape(-1, 4) == sib).all()
        for r in range(len(arity)):
            ev, sib = st["query_round_proofs"][q]["steps"][r]
            a = base + lay.step_evals_offset[r]
            assert (words[a:a + lay.step_evals_words[r]].reshape(-1, 2) == ev).all()
            a = base + lay.step_proof_offset[r]
            assert (words[a:a + 4 * lay.step_siblings[r]].reshape(-1, 4) == sib).all()
    a = lay.final_poly_offset
    assert (words[a:a + lay.final_poly_words].reshape(-1, 2) == st["final_poly"]).all()
    assert int(words[lay.pow_witness_offset]) == st["pow_witness"]
    bad = pkg.FriConfig(rate_bits=2, cap_height=20).fri_params(10)
    with pytest.raises(pkg.BsxError):
        pkg.fri_proof_layout(bad, widths)
