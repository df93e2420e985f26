"""bsx_merkle_reduce_noalloc -- the allocation-free entry point (static device scratch; no counterpart in the reference).
Called through the C ABI and compared bit for bit (integer work: no
tolerance) with the existing bsx_merkle_reduce, on calls in a row that share the scratch: a large reduction, then a
smaller one that must not see the first one's leftovers, then the large one again.  Also device-memory inputs (the leaf
digests of a committed batch reduce to the batch's cap), the size limit and the pool staying untouched.
Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

from conftest import splitmix_field
import __graft_entry__ as G

pytestmark = pytest.mark.gpu
MAX_DIGESTS = 65536  # include/bsx_b200.h BSX_NOALLOC_MAX_DIGESTS


@pytest.fixture(scope="module")
def pkg():
    return G.load_package()


@pytest.fixture(scope="module")
def ctx(pkg):
    c = pkg.Context(0)
    yield c
    c.close()


def _reduce_alloc(pkg, ctx, d, levels):
    out = np.zeros((d.shape[0] >> levels, 4), dtype=np.uint64)
    ctx.check(pkg.lib().bsx_merkle_reduce(ctx._h, d.ctypes.data, d.shape[0], levels, out.ctypes.data))
    return out


def _reduce_noalloc(pkg, ctx, d, levels):
    out = np.zeros((d.shape[0] >> levels, 4), dtype=np.uint64)
    ctx.check(pkg.lib().bsx_merkle_reduce_noalloc(ctx._h, d.ctypes.data, pkg.BSX_HOST, d.shape[0], levels, out.ctypes.data,
                                                  pkg.BSX_HOST))
    return out


def test_noalloc_equals_allocating_entry_point_on_calls_in_a_row(pkg, ctx, oracle):
    big = splitmix_field(9100, (MAX_DIGESTS, 4))
    small = splitmix_field(9101, (64, 4))
    want_big, want_small = _reduce_alloc(pkg, ctx, big, 12), _reduce_alloc(pkg, ctx, small, 6)
    used = ctx.mem_used()
    for _ in range(2):  # big, small, big, small: every call finds the previous call's words in the scratch
        assert (_reduce_noalloc(pkg, ctx, big, 12) == want_big).all()
        assert (_reduce_noalloc(pkg, ctx, small, 6) == want_small).all()
    assert ctx.mem_used() == used  # nothing was taken from (or returned to) the device pool
    # odd and even level counts end in different halves of the scratch; levels = 0 is a copy
    for levels in (0, 1, 2, 5):
        assert (_reduce_noalloc(pkg, ctx, small, levels) == _reduce_alloc(pkg, ctx, small, levels)).all()
    # and the oracle agrees on a case small enough for the CPU
    lvl = [oracle.two_to_one(small[2 * i], small[2 * i + 1]) for i in range(32)]
    assert (_reduce_noalloc(pkg, ctx, small, 1) == np.array(lvl, dtype=np.uint64)).all()


def test_noalloc_device_inputs_reduce_leaf_digests_to_the_cap(pkg, ctx):
    vals = splitmix_field(9102, (9, 1 << 10))
    b = pkg.PolynomialBatch.from_values(ctx, vals, 3, False, 4)
    n_leaves, levels = 1 << 13, 13 - 4
    out = np.zeros((16, 4), dtype=np.uint64)
    for _ in range(2):
        out[:] = 0
        ctx.check(pkg.lib().bsx_merkle_reduce_noalloc(ctx._h, C.c_void_p(pkg.lib().bsx_batch_digests_device(b._h)), pkg.BSX_DEVICE,
                                                      n_leaves, levels, out.ctypes.data, pkg.BSX_HOST))
        assert (out == b.cap).all()
    b.close()


def test_noalloc_rejects_sizes_beyond_its_scratch(pkg, ctx):
    d = np.zeros((2 * MAX_DIGESTS, 4), dtype=np.uint64)
    out = np.zeros((2, 4), dtype=np.uint64)
    rc = pkg.lib().bsx_merkle_reduce_noalloc(ctx._h, d.ctypes.data, pkg.BSX_HOST, d.shape[0], 16, out.ctypes.data, pkg.BSX_HOST)
    assert rc == -4  # BSX_ERR_UNSUPPORTED
    assert b"BSX_NOALLOC_MAX_DIGESTS" in pkg.lib().bsx_last_error(ctx._h)
    rc = pkg.lib().bsx_merkle_reduce_noalloc(ctx._h, d.ctypes.data, pkg.BSX_HOST, 48, 5, out.ctypes.data, pkg.BSX_HOST)
    assert rc == -1  # not a multiple of 2^levels
