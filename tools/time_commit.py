#!/usr/bin/env python3
"""Device-resident commit time without per-phase events (the overlapped LDE / leaf-hash path), CUDA events
on the context's stream.  BSX_NO_OVERLAP=1 forces the sequential path for comparison."""
import argparse, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import __graft_entry__ as G
from bench import splitmix_field
ap = argparse.ArgumentParser()
ap.add_argument("--log-n", type=int, default=20)
ap.add_argument("--cols", type=int, default=135)
ap.add_argument("--rate-bits", type=int, default=3)
ap.add_argument("--reps", type=int, default=5)
a = ap.parse_args()
pkg = G.load_package()
ctx = pkg.Context(0)
stream = torch.cuda.ExternalStream(ctx.stream, device=0)
n = 1 << a.log_n
dev = torch.from_numpy(splitmix_field(0x5EED0001, (a.cols, n)).view(np.int64)).to("cuda:0")
caps = []
for rep in range(a.reps + 2):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    b = pkg.PolynomialBatch.from_values(ctx, None, a.rate_bits, False, 4, device_ptr=dev.data_ptr(), shape=(a.cols, n))
    e1.record(stream)
    torch.cuda.synchronize()
    caps.append(b.cap.copy())
    b.close()
    if rep >= 2:
        print("commit %.2f ms  cap0=%016x" % (e0.elapsed_time(e1), int(caps[-1][0, 0])))
assert all((c == caps[0]).all() for c in caps)
