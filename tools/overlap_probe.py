#!/usr/bin/env python3
"""Experiment: do the ALU-bound transforms and the FMA-heavy-bound leaf hash of DIFFERENT commits overlap when the
hash kernel is kept from filling the SMs (BSX_HASH_SMEM_PAD)?  T threads, one context (stream) each, commit in a
loop; reports aggregate Melem/s."""
import argparse, os, sys, threading, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import __graft_entry__ as G
ap = argparse.ArgumentParser()
ap.add_argument("--threads", type=int, default=2)
ap.add_argument("--reps", type=int, default=4)
ap.add_argument("--log-n", type=int, default=20)
ap.add_argument("--cols", type=int, default=135)
a = ap.parse_args()
pkg = G.load_package()
n = 1 << a.log_n
ctxs = [pkg.Context(0) for _ in range(a.threads)]
devs = [torch.randint(0, 1 << 62, (a.cols, n), dtype=torch.int64, device="cuda:0") for _ in range(a.threads)]
torch.cuda.synchronize()
def work(i, reps):
    for _ in range(reps):
        b = pkg.PolynomialBatch.from_values(ctxs[i], None, 3, False, 4, device_ptr=devs[i].data_ptr(), shape=(a.cols, n))
        b.close()
for i in range(a.threads):
    work(i, 1)
torch.cuda.synchronize()
t0 = time.perf_counter()
th = [threading.Thread(target=work, args=(i, a.reps)) for i in range(a.threads)]
for t in th: t.start()
for t in th: t.join()
torch.cuda.synchronize()
dt = time.perf_counter() - t0
print("pad=%s threads=%d: %.1f ms per commit, %.1f Melem/s aggregate" % (os.environ.get("BSX_HASH_SMEM_PAD", "0"), a.threads,
      dt * 1e3 / (a.threads * a.reps), a.threads * a.reps * n * a.cols / dt / 1e6))
