#!/usr/bin/env python3
"""What pinning a pageable witness in place would cost: cudaHostRegister + cudaHostUnregister of 135 separately allocated
8 MiB vectors (a 2^20 x 135 Vec<PolynomialValues<F>>), against the staged copy of the same bytes (bench.py e2e_pageable).
--procs N runs N processes at once (one per GPU of the box), as eight prover ranks on one host would."""
import argparse, ctypes, os, subprocess, sys, time
import numpy as np
ap = argparse.ArgumentParser()
ap.add_argument("--procs", type=int, default=1)
ap.add_argument("--child", type=int, default=-1)
a = ap.parse_args()
if a.child < 0:
    ps = [subprocess.Popen([sys.executable, os.path.abspath(__file__), "--child", str(i), "--procs", str(a.procs)],
                           stdout=subprocess.PIPE, text=True) for i in range(a.procs)]
    for p in ps:
        print(p.communicate()[0].strip())
    sys.exit(0)
import torch
dev = a.child % max(torch.cuda.device_count(), 1)
torch.cuda.set_device(dev)
torch.zeros(1, device="cuda")
rt = ctypes.CDLL("libcudart.so.12")
vecs = [np.full(1 << 20, i + 1, dtype=np.uint64) for i in range(135)]  # touched: the pages exist
ts = []
for rep in range(4):
    t0 = time.perf_counter()
    for v in vecs:
        rc = rt.cudaHostRegister(ctypes.c_void_p(v.ctypes.data), ctypes.c_size_t(v.nbytes), 0)
        assert rc == 0, rc
    t1 = time.perf_counter()
    for v in vecs:
        assert rt.cudaHostUnregister(ctypes.c_void_p(v.ctypes.data)) == 0
    t2 = time.perf_counter()
    ts.append(((t1 - t0) * 1e3, (t2 - t1) * 1e3))
reg = min(t[0] for t in ts[1:]); unreg = min(t[1] for t in ts[1:])
print("proc %d of %d (gpu %d): register 1.13 GB in 135 pieces %.1f ms, unregister %.1f ms" % (a.child, a.procs, dev, reg, unreg))
