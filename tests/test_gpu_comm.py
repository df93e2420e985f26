"""One process driving G ranks through the C ABI (bsx_comm_create / bsx_sharded_from_values / bsx_sharded_prove_openings):
every word equals the single-GPU batch -- cap, rows, Merkle paths, the whole FRI proof.  The ranks are emulated on
device 0 (`devices = [0] * G`), so this runs on a 1-GPU box; no kernel waits on another rank's kernel."""
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


@pytest.mark.parametrize("world,log_n,n_cols,rate_bits,cap_height", [
    (1, 8, 5, 2, 1), (2, 10, 9, 3, 4), (4, 13, 23, 3, 4), (8, 14, 135, 3, 4), (8, 9, 3, 1, 1), (16, 10, 40, 2, 2),
    (4, 12, 2, 2, 0)])
def test_comm_commit_equals_single_gpu(pkg, ctx, oracle, world, log_n, n_cols, rate_bits, cap_height):
    vals = splitmix_field(3000 + world + log_n, (n_cols, 1 << log_n))
    one = pkg.PolynomialBatch.from_values(ctx, vals, rate_bits, False, cap_height)
    comm = pkg.Comm([0] * world)
    sb = pkg.CommShardedBatch.from_values(comm, vals, rate_bits, cap_height, timing=True)
    assert (sb.cap == one.cap).all()
    nb = 1 << (log_n + rate_bits)
    rng = np.random.default_rng(world)
    leaves = [0, nb - 1] + [int(x) for x in rng.integers(0, nb, 6)]
    rows = sb.get(leaves)
    assert (rows == one.merkle_tree.get(np.array(leaves, dtype=np.uint64))).all()
    for k, leaf in enumerate(leaves):
        path = sb.prove(leaf)
        assert (path == one.merkle_tree.prove(leaf)).all()
        assert oracle.merkle_verify(rows[k], leaf, path, one.cap)
    with pytest.raises(pkg.BsxError):
        sb.get([nb])
    sb.close()
    comm.close()
    one.close()


@pytest.mark.parametrize("world,degree_bits", [(2, 8), (4, 10), (8, 12)])
def test_comm_prove_openings_equals_single_gpu(pkg, ctx, world, degree_bits):
    widths, rate_bits, cap_height = [7, 5, 3], 3, 2
    cfg = pkg.FriConfig(rate_bits=rate_bits, cap_height=cap_height, proof_of_work_bits=6, num_query_rounds=7)
    fp = cfg.fri_params(degree_bits)
    vals = [splitmix_field(4000 + i, (w, 1 << degree_bits)) for i, w in enumerate(widths)]
    single = [pkg.PolynomialBatch.from_values(ctx, v, rate_bits, False, cap_height) for v in vals]
    zeta = splitmix_field(77, (2,))
    g = pow(1753635133440165772, 1 << (32 - degree_bits), pkg.P)
    gz = np.array([int(zeta[0]) * g % pkg.P, int(zeta[1]) * g % pkg.P], dtype=np.uint64)
    polys = [[(o, j) for o, w in enumerate(widths) for j in range(w)], [(2, 0), (2, 1)]]
    inst = pkg.FriInstanceInfo(np.stack([zeta, gz]), polys)
    ch = pkg.Challenger(ctx)
    for b in single:
        ch.observe_cap(b.cap)
    ch2 = ch.copy()
    want = pkg.prove_openings(inst, single, ch, fp)
    comm = pkg.Comm([0] * world)
    sharded = [pkg.CommShardedBatch.from_values(comm, v, rate_bits, cap_height) for v in vals]
    got = pkg.comm_prove_openings(comm, inst, sharded, ch2, fp)
    assert (got == want).all() and ch2.snapshot() == ch.snapshot()
    for b in sharded + single:
        b.close()
    comm.close()
