"""Hand-over from quotient evaluation to the quotient commitment (plonk/prover.rs: `coset_ifft(F::coset_shift())`,
`trim_to_len(quotient_degree)`, `.chunks(degree)`, `PolynomialBatch::from_coeffs`), through the C ABI against the oracle."""
import numpy as np
import pytest

from conftest import splitmix_field
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


@pytest.mark.parametrize("log_n,n_cols", [(0, 1), (1, 3), (5, 2), (10, 9), (13, 8), (16, 3)])
def test_coset_ifft_vs_oracle_and_roundtrip(ctx, oracle, log_n, n_cols):
    v = splitmix_field(900 + log_n, (n_cols, 1 << log_n))
    got = ctx.ntt(v, "coset_ifft", 7)
    for c in range(n_cols):
        assert (got[c] == oracle.coset_ifft(v[c], 7)).all()
    assert (ctx.ntt(got, "coset_fft", 7) == v).all()  # PolynomialCoeffs::coset_fft undoes it


def test_coset_ifft_other_shift_and_bad_shift(pkg, ctx, oracle):
    v = splitmix_field(77, (2, 256))
    shift = 0x123456789ABCDEF
    got = ctx.ntt(v, "coset_ifft", shift)
    assert (got[1] == oracle.coset_ifft(v[1], shift)).all()
    with pytest.raises(pkg.BsxError):
        ctx.ntt(v, "coset_ifft", 0)


@pytest.mark.parametrize("degree_log,q_bits,factor,n_polys,rate_bits,cap_height", [
    (6, 3, 8, 2, 3, 2), (8, 3, 8, 2, 3, 4), (10, 2, 4, 3, 1, 0), (12, 3, 8, 2, 3, 4)])
def test_from_quotient_values_vs_oracle(pkg, ctx, oracle, degree_log, q_bits, factor, n_polys, rate_bits, cap_height):
    vals = splitmix_field(500 + degree_log, (n_polys, 1 << (degree_log + q_bits)))
    b = pkg.PolynomialBatch.from_quotient_values(ctx, vals, degree_log, factor, rate_bits, False, cap_height)
    chunks = oracle.quotient_chunks(vals, degree_log, factor)
    assert chunks.shape == (n_polys * factor, 1 << degree_log)
    want = oracle.batch_from_values(chunks, rate_bits, cap_height, from_coeffs=True)
    assert b.shape()["n_cols"] == n_polys * factor and b.shape()["degree_log"] == degree_log
    assert (b.cap == want["cap"]).all()
    assert (b.polynomials == chunks).all()
    same = pkg.PolynomialBatch.from_coeffs(ctx, chunks, rate_bits, False, cap_height)
    assert (same.cap == b.cap).all()
    for i in (0, 1, (1 << (degree_log + rate_bits)) - 1):
        assert (b.get_lde_values(i, 1) == same.get_lde_values(i, 1)).all()
    b.close()
    same.close()


def test_trimmed_quotient_and_upstream_panic(pkg, ctx, oracle):
    """quotient_degree_factor below 2^q: chunks beyond it must vanish, else upstream's expect() panics."""
    degree_log, q_bits, factor, n = 7, 3, 5, 1 << 7
    coeffs = np.zeros((2, n << q_bits), dtype=np.uint64)
    coeffs[:, :factor * n] = splitmix_field(31, (2, factor * n))
    vals = np.stack([oracle.coset_fft(c, 7) for c in coeffs])
    b = pkg.PolynomialBatch.from_quotient_values(ctx, vals, degree_log, factor, 3, False, 4)
    chunks = oracle.quotient_chunks(vals, degree_log, factor)
    assert (chunks.reshape(2, -1) == coeffs[:, :factor * n]).all()
    assert b.shape()["n_cols"] == 2 * factor and (b.polynomials == chunks).all()
    assert (b.cap == oracle.batch_from_values(chunks, 3, 4, from_coeffs=True)["cap"]).all()
    b.close()
    coeffs[1, factor * n + 3] = 1
    bad = np.stack([oracle.coset_fft(c, 7) for c in coeffs])
    with pytest.raises(ValueError):
        oracle.quotient_chunks(bad, degree_log, factor)
    with pytest.raises(pkg.BsxError):
        pkg.PolynomialBatch.from_quotient_values(ctx, bad, degree_log, factor, 3, False, 4)
    used = ctx.mem_used()
    with pytest.raises(pkg.BsxError):
        pkg.PolynomialBatch.from_quotient_values(ctx, bad, degree_log, 9, 3, False, 4)  # factor > 2^q
    with pytest.raises(pkg.BsxError):
        pkg.PolynomialBatch.from_quotient_values(ctx, bad, degree_log, factor, 3, False, 4)
    assert ctx.mem_used() == used  # failing calls release their device temporaries


@pytest.mark.parametrize("factor", [5, 6, 7])
def test_trim_compaction_does_not_overlap(pkg, ctx, oracle, factor):
    """trim_to_len with keep > big / 2 at a size where the packed and the unpacked ranges of polynomial 1, 2 overlap
    and a copy runs on many CTAs (the in-place cudaMemcpy of round 1 was undefined there)."""
    degree_log, q_bits, n_polys = 16, 3, 3
    n = 1 << degree_log
    coeffs = np.zeros((n_polys, n << q_bits), dtype=np.uint64)
    coeffs[:, :factor * n] = splitmix_field(40 + factor, (n_polys, factor * n))
    vals = ctx.ntt(coeffs, "coset_fft", 7)
    b = pkg.PolynomialBatch.from_quotient_values(ctx, vals, degree_log, factor, 1, False, 4)
    want = coeffs[:, :factor * n].reshape(n_polys * factor, n)
    assert (b.polynomials == want).all()
    same = pkg.PolynomialBatch.from_coeffs(ctx, want, 1, False, 4)
    assert (b.cap == same.cap).all()
    b.close()
    same.close()


def test_from_quotient_values_with_caller_salt(pkg, ctx, oracle):
    """`salt` on the host side (what the Python mirror always passes) must reach the commit as device memory."""
    degree_log, q_bits, factor, rate_bits, cap_height = 8, 3, 8, 2, 3
    vals = splitmix_field(611, (2, 1 << (degree_log + q_bits)))
    salt = splitmix_field(612, (4, 1 << (degree_log + rate_bits)))
    b = pkg.PolynomialBatch.from_quotient_values(ctx, vals, degree_log, factor, rate_bits, True, cap_height, salt=salt)
    chunks = oracle.quotient_chunks(vals, degree_log, factor)
    want = oracle.batch_from_values(chunks, rate_bits, cap_height, salt=salt, from_coeffs=True)
    assert (b.cap == want["cap"]).all()
    same = pkg.PolynomialBatch.from_coeffs(ctx, chunks, rate_bits, True, cap_height, salt=salt)
    assert (same.cap == b.cap).all()
    b.close()
    same.close()
