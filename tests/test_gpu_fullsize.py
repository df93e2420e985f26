"""Oracle parity at the sizes the bench and the sweep report numbers for: the headline
2^20 x 135 commit (rate_bits 3, cap_height 4) and the sweep corners that take other transform plans.  One CPU
commit each (tens of seconds on the box's host cores); the cap is compared bit for bit, and at the headline also
three LDE rows and three Merkle paths.  Parity is against the in-repo restatement of plonky2 (parity unpinned
against the real prover: the reference mount is empty)."""
import numpy as np
import pytest

from conftest import splitmix_field
import __graft_entry__ as G

pytestmark = [pytest.mark.gpu, pytest.mark.slow]


@pytest.fixture(scope="module")
def pkg():
    return G.load_package()


@pytest.fixture(scope="module")
def ctx(pkg):
    c = pkg.Context(0)
    yield c
    c.close()


def _bitrev(x, bits):
    return int(format(x, "0%db" % bits)[::-1], 2) if bits else 0


def test_headline_commit_vs_oracle(pkg, ctx, oracle):
    log_n, n_cols, rate_bits, cap_height = 20, 135, 3, 4
    log_nb = log_n + rate_bits
    vals = splitmix_field(0x5EED0001, (n_cols, 1 << log_n))
    b = pkg.PolynomialBatch.from_values(ctx, vals, rate_bits, False, cap_height)
    cap = b.cap
    leaf_idx = [0, 0x2AAAAA % (1 << log_nb), (1 << log_nb) - 1]
    rows = b.merkle_tree.get(np.array(leaf_idx, dtype=np.uint64))
    paths = [b.merkle_tree.prove(i) for i in leaf_idx]
    lde_rows = [b.get_lde_values(_bitrev(i, log_nb), 1) for i in leaf_idx]
    b.close()
    want = oracle.batch_from_values(vals, rate_bits, cap_height, want=("leaves", "digests"))
    assert (cap == want["cap"]).all()
    for k, i in enumerate(leaf_idx):
        assert (rows[k] == want["leaves"][i]).all()
        assert (lde_rows[k] == want["leaves"][i]).all()
        sib = oracle.merkle_prove(want["digests"], log_nb, cap_height, i)
        assert (paths[k] == sib).all()
        assert oracle.merkle_verify(rows[k], i, paths[k], cap)


@pytest.mark.parametrize("log_n,n_cols,rate_bits", [(16, 80, 1), (18, 400, 4), (22, 80, 1)])
def test_sweep_corner_caps_vs_oracle(pkg, ctx, oracle, log_n, n_cols, rate_bits):
    vals = splitmix_field(0x5EED0005 + log_n, (n_cols, 1 << log_n))
    b = pkg.PolynomialBatch.from_values(ctx, vals, rate_bits, False, 4)
    cap = b.cap
    b.close()
    want = oracle.batch_from_values(vals, rate_bits, 4, want=())
    assert (cap == want["cap"]).all()
