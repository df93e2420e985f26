#!/usr/bin/env python3
"""Time `Context.partial_products_and_zs` on device-resident wires / sigmas: 80 routed wires x 2^20 rows, two challenges,
max_degree 8 (standard_recursion_config) -> the 20 columns of the Zs commitment."""
import argparse, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import __graft_entry__ as G
ap = argparse.ArgumentParser()
ap.add_argument("--log-n", type=int, default=20)
ap.add_argument("--routed", type=int, default=80)
ap.add_argument("--degree", type=int, default=8)
ap.add_argument("--challenges", type=int, default=2)
ap.add_argument("--reps", type=int, default=5)
a = ap.parse_args()
pkg = G.load_package()
ctx = pkg.Context(0)
stream = torch.cuda.ExternalStream(ctx.stream, device=0)
n = 1 << a.log_n
gen = torch.Generator(device="cuda:0"); gen.manual_seed(4)
w = torch.randint(0, 1 << 62, (a.routed, n), dtype=torch.int64, device="cuda:0", generator=gen)
s = torch.randint(0, 1 << 62, (a.routed, n), dtype=torch.int64, device="cuda:0", generator=gen)
n_chunks = -(-a.routed // a.degree)
out = torch.empty((a.challenges * n_chunks, n), dtype=torch.int64, device="cuda:0")
k_is = np.array([pow(7, j, pkg.P) for j in range(a.routed)], dtype=np.uint64)
betas = np.arange(3, 3 + a.challenges, dtype=np.uint64)
gammas = np.arange(9, 9 + a.challenges, dtype=np.uint64)
torch.cuda.synchronize()
ts = []
for rep in range(a.reps + 2):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    ctx.partial_products_and_zs(None, None, k_is, betas, gammas, a.degree, out_device_ptr=out.data_ptr(),
                                device_ptrs=(w.data_ptr(), s.data_ptr()), shape=(a.routed, n))
    e1.record(stream)
    torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1))
rd = 2 * a.routed * n * 8 * a.challenges
print("partial_products_and_zs %d routed x 2^%d, %d challenges, max_degree %d: %.2f ms (best %.2f), %.0f GB/s of wire/sigma reads"
      % (a.routed, a.log_n, a.challenges, a.degree, float(np.median(ts[2:])), min(ts[2:]), rd / (min(ts[2:]) * 1e-3) / 1e9))
