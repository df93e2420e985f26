"""CPU tests of the row-sharded commit's host logic: the ownership arithmetic of
`ShardPlan` against the C ABI's own `bsx_shard_column_range`, the staging-layout contract, the cap
assembly, and the all-to-all exchange itself over `gloo` with world sizes 2 and 4 (no GPU involved: the
LDE words are synthetic, the layout is what is tested)."""
import ctypes as C
import os
import socket
import sys

import numpy as np
import pytest

from conftest import ROOT
import __graft_entry__ as G


@pytest.fixture(scope="module")
def pkg():
    return G.load_package()


@pytest.mark.parametrize("n_cols,world", [(135, 8), (135, 4), (135, 2), (135, 1), (80, 8), (3, 8), (0, 2), (400, 16)])
def test_column_ranges_match_the_abi(pkg, n_cols, world):
    plan = pkg.ShardPlan(10, n_cols, 3, 4, world)
    L = pkg.lib()
    covered = []
    for r in range(world):
        c0, c1 = C.c_int32(), C.c_int32()
        assert L.bsx_shard_column_range(n_cols, r, world, C.byref(c0), C.byref(c1)) == 0
        assert (c0.value, c1.value) == plan.column_range(r)
        covered += list(range(c0.value, c1.value))
    assert covered == list(range(n_cols))
    sizes = [plan.column_range(r)[1] - plan.column_range(r)[0] for r in range(world)]
    assert max(sizes) - min(sizes) <= 1 and sizes == sorted(sizes, reverse=True)
    assert L.bsx_shard_column_range(n_cols, world, world, C.byref(c0), C.byref(c1)) != 0


def test_plan_ownership(pkg):
    plan = pkg.ShardPlan(20, 135, 3, 4, 8)
    assert plan.chunk == 1 << 20 and plan.local_cap_height == 1 and plan.reduce_levels == 0
    assert plan.owner(0) == 0 and plan.owner((1 << 23) - 1) == 7 and plan.owner(3 << 20) == 3
    assert plan.local_leaf((3 << 20) + 5) == 5
    # headline exchange volume: each rank sends 7/8 of its LDE columns
    sent = sum(plan.staging_block_words(0) for d in range(1, 8)) * 8
    assert sent == 17 * (1 << 20) * 7 * 8
    deep = pkg.ShardPlan(4, 5, 1, 1, 8)  # more ranks than cap entries: rank roots are hashed up
    assert deep.local_cap_height == 0 and deep.reduce_levels == 2
    with pytest.raises(pkg.BsxError):
        pkg.ShardPlan(1, 5, 1, 0, 8)
    with pytest.raises(pkg.BsxError):
        pkg.ShardPlan(4, 5, 1, 0, 3)
    with pytest.raises(pkg.BsxError):
        pkg.ShardPlan(4, 5, 1, 6, 2)


def test_assemble_cap_concatenates_in_rank_order(pkg):
    plan = pkg.ShardPlan(6, 7, 2, 3, 4)
    caps = [np.arange(8, dtype=np.uint64).reshape(2, 4) + 100 * r for r in range(4)]
    cap = pkg.assemble_cap(caps, plan)
    assert cap.shape == (8, 4) and (cap[2] == caps[1][0]).all() and (cap[7] == caps[3][1]).all()
    with pytest.raises(pkg.BsxError):  # hashing rank roots needs the device
        pkg.assemble_cap([c[:1] for c in caps], pkg.ShardPlan(6, 7, 2, 1, 4))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, log_n, n_cols, rate_bits, out):
    import torch
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    import __graft_entry__ as G2
    pkg = G2.load_package()
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        plan = pkg.ShardPlan(log_n, n_cols, rate_bits, 1, world)
        nb = 1 << plan.log_nb
        # synthetic "LDE": word (col, leaf) = col * 2^32 + leaf
        c0, c1 = plan.column_range(rank)
        send = np.zeros(world * plan.staging_block_words(rank), dtype=np.int64)
        for c in range(c0, c1):
            for leaf in range(nb):
                send[plan.staging_index(rank, c, leaf)] = (c << 32) + leaf
        recv = torch.full((n_cols * plan.chunk,), -1, dtype=torch.int64)
        pkg.exchange_staged(torch.from_numpy(send), recv, plan, rank)
        slab = recv.numpy().reshape(n_cols, plan.chunk)
        want = (np.arange(n_cols, dtype=np.int64)[:, None] << 32) + (rank * plan.chunk + np.arange(plan.chunk))[None, :]
        ok = bool((slab == want).all())
        # the cap all-gather of from_values, on the same group
        mine = torch.arange(8, dtype=torch.int64) + 1000 * rank
        allc = [torch.empty(8, dtype=torch.int64) for _ in range(world)]
        dist.all_gather(allc, mine)
        cap = pkg.assemble_cap([a.numpy().astype(np.uint64) for a in allc], pkg.ShardPlan(log_n, n_cols, rate_bits,
                                                                                       1 + plan.log_world, world))
        ok = ok and cap.shape == (2 * world, 4) and int(cap[2 * rank, 0]) == 1000 * rank
        out.put((rank, ok))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,n_cols", [(2, 5), (4, 7)])
def test_exchange_layout_over_gloo(world, n_cols):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, 5, n_cols, 2, out)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted(out.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert res == [(r, True) for r in range(world)]
