"""The row-sharded commit against the single-GPU batch and the CPU oracle.  G ranks are emulated on ONE
device (G shards in one process, slabs wired with bsx_shard_set_peer_ptr, or the staged all-to-all done
with device copies), so the whole sharded data path -- column split, NTT store redirection, per-rank
leaf hashing and cap subtrees, cap assembly -- runs on the 1-GPU box.  The multi-process version (CUDA IPC
+ NCCL, one process per GPU) is tests/mp_shard_worker.py, run here when the box has >= 2 GPUs."""
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import P, ROOT, splitmix_field
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


def _sharded(pkg, ctx, vals, rate_bits, cap_height, world, mode, is_coeffs=False):
    import torch
    n_cols, n = vals.shape
    plan = pkg.ShardPlan(n.bit_length() - 1, n_cols, rate_bits, cap_height, world)
    shards = [pkg.ShardedPolynomialBatch.create(ctx, plan, r) for r in range(world)]
    if mode == "p2p":
        for a in shards:
            for b in shards:
                if a is not b:
                    a.set_peer_ptr(b.rank, b.slab_ptr)
    for s in shards:
        c0, c1 = plan.column_range(s.rank)
        s.transform(vals[c0:c1], is_coeffs=is_coeffs, use_staging=(mode == "staged"))
    if mode == "staged":  # what exchange_staged does over NCCL, with device copies
        for s in shards:
            blk = plan.staging_block_words(s.rank)
            send = pkg.device_tensor(s.staging_ptr, world * blk, 0)
            for d in shards:
                recv = pkg.device_tensor(d.slab_ptr, n_cols * plan.chunk, 0)
                off = plan.slab_offset_words(s.rank)
                recv[off:off + blk].copy_(send[d.rank * blk:(d.rank + 1) * blk])
        torch.cuda.synchronize()
    caps = [s.commit() for s in shards]
    return plan, shards, pkg.assemble_cap(caps, plan, ctx)


@pytest.mark.parametrize("log_n,n_cols,rate_bits,cap_height,world,mode", [
    (10, 9, 3, 4, 2, "p2p"), (10, 9, 3, 4, 8, "p2p"), (13, 19, 3, 4, 4, "p2p"), (14, 135, 3, 4, 8, "p2p"),
    (10, 9, 3, 4, 4, "staged"), (13, 19, 1, 4, 8, "staged"), (14, 135, 3, 4, 8, "staged"),
    (6, 5, 2, 0, 4, "p2p"), (5, 3, 1, 1, 8, "staged"), (12, 8, 2, 3, 1, "p2p"), (3, 2, 0, 1, 8, "p2p"),
])
def test_sharded_equals_single_gpu(pkg, ctx, oracle, log_n, n_cols, rate_bits, cap_height, world, mode):
    vals = splitmix_field(7000 + log_n + world, (n_cols, 1 << log_n))
    vals[0, :4] = np.array([0, P - 1, 2**32 - 1, 2**32], dtype=np.uint64)[:min(4, 1 << log_n)]
    ref = pkg.PolynomialBatch.from_values(ctx, vals, rate_bits, False, cap_height)
    plan, shards, cap = _sharded(pkg, ctx, vals, rate_bits, cap_height, world, mode)
    assert (cap == ref.cap).all(), "sharded cap differs from the single-GPU cap"
    want = oracle.batch_from_values(vals, rate_bits, cap_height, want=())
    assert (cap == want["cap"]).all(), "sharded cap differs from the CPU oracle"
    # coefficients: each rank holds its columns
    polys = ref.polynomials
    for s in shards:
        c0, c1 = plan.column_range(s.rank)
        assert (s.polynomials == polys[c0:c1]).all()
    # openings: rows and Merkle paths served by the owning rank equal the single-GPU tree's
    nb = 1 << plan.log_nb
    tree = ref.merkle_tree
    for leaf in sorted({0, 1, nb - 1, nb // 2, (nb // 3) | 1, plan.chunk - 1, plan.chunk % nb}):
        s = shards[plan.owner(leaf)]
        assert (s.merkle_tree.get(plan.local_leaf(leaf)) == tree.get(leaf)).all()
        if plan.reduce_levels == 0:
            assert (s.merkle_tree.prove(plan.local_leaf(leaf)) == tree.prove(leaf)).all()
    # the same openings as the collective `open` assembles them: every rank contributes the leaves it owns,
    # with the proof extended to the cap of the whole batch when there are more ranks than cap entries
    leaves = sorted({0, 1, nb - 1, nb // 2, (nb // 3) | 1, plan.chunk - 1, plan.chunk % nb})
    caps = [s.merkle_tree.cap for s in shards]
    rows = proofs = 0
    for s in shards:
        s.rank_caps = caps
        r, pr = s.local_openings(leaves)
        rows, proofs = rows + r, proofs + pr
    for k, leaf in enumerate(leaves):
        assert (rows[k] == tree.get(leaf)).all()
        assert (proofs[k] == np.asarray(tree.prove(leaf)).reshape(-1, 4)).all()
        assert oracle.merkle_verify(rows[k], leaf, proofs[k], cap)
        with pytest.raises(pkg.BsxError):
            shards[(plan.owner(leaf) + 1) % world].get(leaf) if world > 1 else shards[0].get(nb)
    for s in shards:
        s.close()
    ref.close()


def test_sharded_from_coeffs(pkg, ctx):
    co = splitmix_field(99, (11, 1 << 13))
    ref = pkg.PolynomialBatch.from_coeffs(ctx, co, 3, False, 4)
    plan, shards, cap = _sharded(pkg, ctx, co, 3, 4, 4, "p2p", is_coeffs=True)
    assert (cap == ref.cap).all()
    for s in shards:
        s.close()
    ref.close()


def test_shard_errors(pkg, ctx):
    with pytest.raises(pkg.BsxError):
        pkg.ShardedPolynomialBatch.create(ctx, pkg.ShardPlan(4, 3, 1, 9, 2), 0)  # cap_height > log2(leaves)
    plan = pkg.ShardPlan(6, 4, 1, 1, 2)
    s = pkg.ShardedPolynomialBatch.create(ctx, plan, 0)
    with pytest.raises(pkg.BsxError):  # peer slab never wired
        s.transform(splitmix_field(1, (2, 64)))
    with pytest.raises(pkg.BsxError):  # wrong column count for this rank
        s.transform(splitmix_field(1, (3, 64)), use_staging=True)
    s.close()


def test_merkle_reduce(pkg, ctx, oracle):
    d = splitmix_field(5, (8, 4))
    out = np.zeros((2, 4), dtype=np.uint64)
    ctx.check(pkg.lib().bsx_merkle_reduce(ctx._h, d.ctypes.data, 8, 2, out.ctypes.data))
    lvl = [oracle.two_to_one(d[2 * i], d[2 * i + 1]) for i in range(4)]
    want = [oracle.two_to_one(lvl[0], lvl[1]), oracle.two_to_one(lvl[2], lvl[3])]
    assert (out == np.array(want, dtype=np.uint64)).all()


@pytest.mark.parametrize("mode", ["p2p", "p2p_host", "nccl"])
def test_multiprocess_ipc_and_nccl(mode):
    import torch
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs (one process per GPU, CUDA IPC / NCCL)")
    world = 1 << (min(n, 8).bit_length() - 1)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world),
           "--master-addr", "127.0.0.1", "--master-port", "29731", os.path.join(ROOT, "tests", "mp_shard_worker.py"),
           "--mode", mode]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    assert "sharded commit ok" in r.stdout


@pytest.mark.parametrize("world,cap_height", [(2, 2), (4, 1), (8, 4)])
def test_sharded_prove_openings_equals_single_gpu(pkg, ctx, world, cap_height):
    """prove_openings over row-sharded oracles (ranks emulated on one device): parts of the final polynomial from the
    ranks' coefficient columns, summed; FRI from the final polynomial; initial-tree openings served by the owners and
    spliced in.  Word for word the single-GPU proof, including world > 2^cap_height (paths extended above the rank
    roots)."""
    log_n, rate_bits, widths = 10, 2, (9, 5)
    vals = [splitmix_field(7300 + i, (w, 1 << log_n)) for i, w in enumerate(widths)]
    fresh = [pkg.PolynomialBatch.from_values(ctx, v, rate_bits, False, cap_height) for v in vals]
    sets = []  # sets[o][rank]
    for v in vals:
        plan, shards, cap = _sharded(pkg, ctx, v, rate_bits, cap_height, world, "p2p")
        caps = [s.merkle_tree.cap for s in shards]
        for s in shards:
            s.rank_caps, s.cap = caps, cap
        sets.append(shards)
    zeta = splitmix_field(7310, (2,))
    w = pow(7, (P - 1) >> log_n, P)
    gz = np.array([int(zeta[0]) * w % P, int(zeta[1]) * w % P], dtype=np.uint64)
    polys = [[(o, j) for o, wd in enumerate(widths) for j in range(wd)], [(0, 1), (1, 4)]]
    inst = pkg.FriInstanceInfo(np.stack([zeta, gz]), polys)
    fp = pkg.FriConfig(rate_bits=rate_bits, cap_height=cap_height, proof_of_work_bits=5, num_query_rounds=6).fri_params(log_n)

    def transcript():
        ch = pkg.Challenger(ctx)
        for b in fresh:
            ch.observe_cap(b.cap)
        return ch

    want = pkg.prove_openings(inst, fresh, transcript(), fp)
    # every rank's part, with the challenger every rank would hold
    alpha = transcript().get_extension_challenge()
    parts = [pkg.final_poly_part([sets[o][r] for o in range(len(widths))], inst, alpha) for r in range(world)]
    assert (pkg.sum_parts(ctx, parts) == pkg.final_poly(inst, fresh, alpha)).all()

    def opener(sb, leaves):  # what the all-reduce in ShardedPolynomialBatch.open adds up
        o = [k for k in range(len(widths)) if sb in sets[k]][0]
        rows = proofs = 0
        for s in sets[o]:
            r, pr = s.local_openings(leaves)
            rows, proofs = rows + r, proofs + pr
        return rows, proofs

    for rank in (0, world - 1):
        mine = [sets[o][rank] for o in range(len(widths))]
        ch = transcript()
        got = pkg.sharded_prove_openings(inst, mine, ch, fp, gather=lambda part: parts, opener=opener)
        assert got.shape == want.shape and (got == want).all()
    for shards in sets:
        for s in shards:
            s.close()
    for b in fresh:
        b.close()


def test_device_barrier_refuses_two_processes_on_one_gpu(tmp_path):
    """bsx_shard_run_p2p spins on flags that the peer's kernel writes: two ranks on ONE GPU (here two processes with
    IPC-mapped slabs) could wait on each other, so the call must refuse them (device tag behind the mailbox) before it
    launches anything."""
    worker = os.path.join(ROOT, "tests", "ipc_same_device_worker.py")
    ps = [subprocess.Popen([sys.executable, worker, str(r), str(tmp_path)], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
          for r in range(2)]
    outs = [p.communicate(timeout=180) for p in ps]
    for p, (out, err) in zip(ps, outs):
        assert p.returncode == 0, err[-2000:]
        assert "rc=-4" in out and "shares device" in out, out + err[-1000:]
