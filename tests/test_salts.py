"""Blinding salts (fri/oracle.rs lde_values `F::rand_vec`): ChaCha20 key stream under a per-batch key from getrandom."""
import numpy as np
import pytest

import chacha_ref as R
import __graft_entry__ as G

P = 0xFFFFFFFF00000001


def test_python_chacha_matches_rfc8439_vector():
    assert R.chacha20_block_words(R.RFC8439_KEY, 1, 0x09000000, 0x4A000000, 0) == R.RFC8439_BLOCK


@pytest.fixture(scope="module")
def pkg():
    return G.load_package()


@pytest.fixture(scope="module")
def ctx(pkg):
    c = pkg.Context(0)
    yield c
    c.close()


@pytest.mark.gpu
def test_device_stream_is_chacha20(ctx):
    for key in (R.RFC8439_KEY, [0] * 8, [0xFFFFFFFF] * 8):
        got = ctx.salt_stream(key, 8 * 300 + 5, raw=True)
        want = np.array(R.salt_stream_raw(key, got.size), dtype=np.uint64)
        assert (got == want).all()
        salts = ctx.salt_stream(key, 8 * 300 + 5)
        assert (salts < np.uint64(P)).all()
        ok = want < np.uint64(P)
        assert (salts[ok] == want[ok]).all()  # accepted candidates pass through unchanged


@pytest.mark.gpu
def test_two_blinded_batches_get_unrelated_salts(pkg, ctx):
    vals = np.arange(3 * 64, dtype=np.uint64).reshape(3, 64)
    a = pkg.PolynomialBatch.from_values(ctx, vals, 1, True, 2)
    b = pkg.PolynomialBatch.from_values(ctx, vals, 1, True, 2)
    la, lb = a.export()["leaves"], b.export()["leaves"]
    assert la.shape == (128, 3 + 4)
    assert (la[:, :3] == lb[:, :3]).all()
    sa, sb = la[:, 3:].reshape(-1), lb[:, 3:].reshape(-1)
    assert (sa < np.uint64(P)).all() and (sb < np.uint64(P)).all()
    assert len(set(sa.tolist())) == sa.size  # 512 words of a 64-bit stream: no repeats
    assert not (sa == sb).any() and not (a.cap == b.cap).all()
    a.close()
    b.close()
