"""torchrun worker: one process per GPU commits one batch with ShardedPolynomialBatch.from_values and
checks cap, rows and Merkle paths against the single-GPU batch computed on rank 0's device."""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__ as G  # noqa: E402
from conftest import splitmix_field  # noqa: E402


def main():
    import torch
    import torch.distributed as dist
    ap = argparse.ArgumentParser()
    ap.add_argument("--mode", default="p2p")
    ap.add_argument("--log-n", type=int, default=14)
    ap.add_argument("--cols", type=int, default=135)
    a = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    pkg = G.load_package()
    ctx = pkg.Context(local)
    vals = splitmix_field(4242, (a.cols, 1 << a.log_n))  # every rank generates the same witness
    plan = pkg.ShardPlan(a.log_n, a.cols, 3, 4, world)
    c0, c1 = plan.column_range(rank)
    sb = pkg.ShardedPolynomialBatch.from_values(ctx, vals[c0:c1], a.log_n, a.cols, 3, 4, mode=a.mode)
    ref = pkg.PolynomialBatch.from_values(ctx, vals, 3, False, 4)
    assert (sb.cap == ref.cap).all(), "rank %d: sharded cap differs" % rank
    tree = ref.merkle_tree
    for k in range(16):
        leaf = rank * plan.chunk + (k * 2654435761) % plan.chunk
        assert (sb.merkle_tree.get(plan.local_leaf(leaf)) == tree.get(leaf)).all()
        assert (sb.merkle_tree.prove(plan.local_leaf(leaf)) == tree.prove(leaf)).all()
    assert (sb.polynomials == ref.polynomials[c0:c1]).all()
    # repeated commits on the same slabs (epochs of the device-side barrier), once from pinned host memory (the chunked
    # upload under the transforms) and once from device memory
    pin = torch.from_numpy(vals[c0:c1].view(np.int64).copy()).pin_memory()
    sb.run(pin.numpy().view(np.uint64))
    assert (sb.cap == ref.cap).all(), "rank %d: second commit (pinned host input) differs" % rank
    devt = pin.to("cuda:%d" % local)
    torch.cuda.synchronize()
    sb.run(None, device_ptr=devt.data_ptr())
    assert (sb.cap == ref.cap).all(), "rank %d: third commit (device input) differs" % rank
    # collective openings: every rank asks for the same leaves, each answered by its owner
    nb = 1 << plan.log_nb
    leaves = [(k * 40503 + 17) % nb for k in range(28)]
    rows, proofs = sb.open(leaves)
    for k, leaf in enumerate(leaves):
        assert (rows[k] == tree.get(leaf)).all()
        assert (proofs[k] == np.asarray(tree.prove(leaf)).reshape(-1, 4)).all()
    # prove_openings over the sharded oracle: every rank ends with the single-GPU proof
    zeta = splitmix_field(4243, (1, 2))
    inst = pkg.FriInstanceInfo(zeta, [[(0, j) for j in range(a.cols)]])
    fp = pkg.FriConfig(rate_bits=3, cap_height=4, proof_of_work_bits=8, num_query_rounds=10).fri_params(a.log_n)
    ch_s, ch_r = pkg.Challenger(ctx), pkg.Challenger(ctx)
    ch_s.observe_cap(sb.cap)
    ch_r.observe_cap(ref.cap)
    proof_s = pkg.sharded_prove_openings(inst, [sb], ch_s, fp)
    proof_r = pkg.prove_openings(inst, [ref], ch_r, fp)
    assert (proof_s == proof_r).all(), "rank %d: sharded proof differs" % rank
    assert ch_s.snapshot() == ch_r.snapshot()
    cap0 = int(sb.cap[0, 0])
    dist.barrier()
    sb.close()
    ref.close()
    ctx.close()
    if rank == 0:
        print("sharded commit ok: mode=%s world=%d cap0=%016x" % (a.mode, world, cap0))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
