"""Opening-set evaluation: `eval_commitment` on the device vs the oracle's Horner loop."""
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


@pytest.mark.parametrize("degree_bits", [0, 1, 5, 8, 11, 12, 13, 15])
@pytest.mark.parametrize("from_coeffs", [False, True])
def test_eval_commitment_vs_oracle(pkg, ctx, oracle, degree_bits, from_coeffs):
    n_cols = 5
    vals = splitmix_field(70 + degree_bits, (n_cols, 1 << degree_bits))
    ob = oracle.batch_from_values(vals, 1, 0, from_coeffs=from_coeffs)
    mk = pkg.PolynomialBatch.from_coeffs if from_coeffs else pkg.PolynomialBatch.from_values
    gb = mk(ctx, vals, 1, False, 0)
    pts = splitmix_field(700 + degree_bits, (3, 2))
    pts[2, 1] = 0  # a base-field point
    got = gb.eval_commitment(pts)
    assert got.shape == (3, n_cols, 2)
    for k in range(3):
        for j in range(n_cols):
            want = oracle.eval_poly_ext(ob, j, pts[k], 1, 0)
            assert (got[k, j] == want).all(), (k, j)
    # single point, column sub-range
    one = gb.eval_commitment(pts[0], col_begin=1, n_cols=3)
    assert (one == got[0, 1:4]).all()
    gb.close()


def test_eval_edge_values_and_errors(pkg, ctx, oracle):
    vals = np.full((2, 64), P - 1, dtype=np.uint64)
    vals[1, :] = 0
    gb = pkg.PolynomialBatch.from_coeffs(ctx, vals, 1, False, 0)
    ob = oracle.batch_from_values(vals, 1, 0, from_coeffs=True)
    z = np.array([P - 1, P - 1], dtype=np.uint64)
    got = gb.eval_commitment(z)
    assert (got[0] == oracle.eval_poly_ext(ob, 0, z, 1, 0)).all()
    assert (got[1] == 0).all()
    assert gb.eval_commitment(np.zeros((0, 2), np.uint64)).shape == (0, 2, 2)
    with pytest.raises(pkg.BsxError):
        gb.eval_commitment(z, col_begin=1, n_cols=2)
    gb.close()


@pytest.mark.parametrize("degree_bits,n_cols", [(20, 135), (23, 3)])
def test_eval_on_the_coset_equals_lde_row(pkg, ctx, degree_bits, n_cols):
    """Size-independent tie to the commit: p(g * w^i) is row i of the LDE the Merkle tree was built on."""
    rate_bits = 1
    vals = splitmix_field(91, (n_cols, 1 << degree_bits))
    gb = pkg.PolynomialBatch.from_values(ctx, vals, rate_bits, False, 4)
    nb_log = degree_bits + rate_bits
    w = pow(7, (P - 1) >> nb_log, P)  # generator 7: primitive root of the 2^nb_log subgroup
    idx = [0, 1, 12345, (1 << nb_log) - 1]
    pts = np.array([[7 * pow(w, i, P) % P, 0] for i in idx], dtype=np.uint64)
    got = gb.eval_commitment(pts)
    for k, i in enumerate(idx):
        row = gb.get_lde_values(i)
        assert (got[k, :, 0] == row).all() and (got[k, :, 1] == 0).all()
    gb.close()
