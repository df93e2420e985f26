"""Tier 3: property tests through the C ABI, `hypothesis` drawing shapes and values.
Bit-exact against the oracle; adversarial words (0, 1, p-1, 2^32-1, 2^32, p - 2^32) are mixed into every draw so
the reduction carry/borrow paths are hit."""
import numpy as np
import pytest
from hypothesis import HealthCheck, given, settings, strategies as st

from conftest import P, splitmix_field
import __graft_entry__ as G

pytestmark = pytest.mark.gpu

EDGE = np.array([0, 1, P - 1, 2**32 - 1, 2**32, P - 2**32, 2**63, P - 2], dtype=np.uint64)
SETTINGS = dict(deadline=None, suppress_health_check=list(HealthCheck), derandomize=True)


@pytest.fixture(scope="module")
def pkg():
    return G.load_package()


@pytest.fixture(scope="module")
def ctx(pkg):
    c = pkg.Context(0)
    yield c
    c.close()


def _field(seed, shape, edge_every):
    a = splitmix_field(seed, shape).reshape(-1)
    if edge_every and a.size:
        idx = np.arange(0, a.size, edge_every)
        a[idx] = EDGE[(idx // edge_every + seed) % EDGE.size]
    return a.reshape(shape)


@settings(max_examples=40, **SETTINGS)
@given(log_n=st.integers(0, 11), n_cols=st.integers(1, 24), rate_bits=st.integers(0, 4), cap_frac=st.floats(0, 1),
       blinding=st.booleans(), from_coeffs=st.booleans(), seed=st.integers(0, 2**31), edge_every=st.sampled_from([0, 1, 3, 17]))
def test_commit_matches_oracle(pkg, ctx, oracle, log_n, n_cols, rate_bits, cap_frac, blinding, from_coeffs, seed, edge_every):
    log_nb = log_n + rate_bits
    cap_height = int(round(cap_frac * log_nb))
    vals = _field(seed, (n_cols, 1 << log_n), edge_every)
    salt = _field(seed + 1, (4, 1 << log_nb), edge_every) if blinding else None
    mk = pkg.PolynomialBatch.from_coeffs if from_coeffs else pkg.PolynomialBatch.from_values
    b = mk(ctx, vals, rate_bits, blinding, cap_height, salt=salt)
    want = oracle.batch_from_values(vals, rate_bits, cap_height, salt=salt, from_coeffs=from_coeffs)
    assert (b.cap == want["cap"]).all()
    ex = b.export()
    assert (ex["coeffs"] == want["coeffs"]).all() and (ex["leaves"] == want["leaves"]).all()
    if want["digests"] is not None and want["digests"].size:
        assert (ex["digests"].reshape(-1) ==
                oracle.merkle_digests_to_plonky2(want["digests"], log_nb, cap_height).reshape(-1)).all()
    leaf = seed % (1 << log_nb)
    tree = b.merkle_tree
    assert oracle.merkle_verify(tree.get(leaf), leaf, tree.prove(leaf), b.cap)
    z = _field(seed + 2, (2,), 0)
    got = b.eval_commitment(z)
    j = seed % n_cols
    assert (got[j] == oracle.eval_poly_ext(want, j, z, rate_bits, cap_height)).all()
    b.close()


@settings(max_examples=30, **SETTINGS)
@given(n=st.integers(1, 700), seed=st.integers(0, 2**31), edge_every=st.sampled_from([0, 1, 2, 5]))
def test_poseidon_matches_oracle(ctx, oracle, n, seed, edge_every):
    states = _field(seed, (n, 12), edge_every)
    assert (ctx.poseidon_permute(states.copy()) == oracle.poseidon(states.copy())).all()


@settings(max_examples=30, **SETTINGS)
@given(log_n=st.integers(0, 14), n_cols=st.integers(1, 12), kind=st.sampled_from(["fft", "ifft", "coset_fft"]),
       seed=st.integers(0, 2**31), edge_every=st.sampled_from([0, 1, 7]))
def test_ntt_matches_oracle(ctx, oracle, log_n, n_cols, kind, seed, edge_every):
    cols = _field(seed, (n_cols, 1 << log_n), edge_every)
    got = ctx.ntt(cols.copy(), kind)
    fn = {"fft": oracle.fft, "ifft": oracle.ifft, "coset_fft": oracle.coset_fft}[kind]
    for c in range(n_cols):
        assert (got[c] == fn(cols[c].copy())).all()


@settings(max_examples=25, **SETTINGS)
@given(log_leaves=st.integers(0, 10), leaf_len=st.integers(1, 40), cap_frac=st.floats(0, 1), seed=st.integers(0, 2**31))
def test_merkle_tree_matches_oracle(pkg, ctx, oracle, log_leaves, leaf_len, cap_frac, seed):
    cap_height = int(round(cap_frac * log_leaves))
    leaves = _field(seed, (1 << log_leaves, leaf_len), 5)
    t = pkg.MerkleTree(ctx, leaves, cap_height)
    digests, cap = oracle.merkle_build(leaves, cap_height)
    assert (t.cap == cap).all()
    i = seed % (1 << log_leaves)
    assert (np.asarray(t.prove(i)).reshape(-1, 4) ==
            np.asarray(oracle.merkle_prove(digests, log_leaves, cap_height, i)).reshape(-1, 4)).all()
    t.close()
