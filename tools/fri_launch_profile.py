#!/usr/bin/env python3
"""One prove_openings at 2^20 (255 polynomials in 4 batches) for an ncu launch list: the commits happen first, the
profiler range covers prove_openings only (cudaProfilerStart/Stop; run ncu with --profile-from-start off)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import __graft_entry__ as G
from bench import splitmix_field
pkg = G.load_package()
ctx = pkg.Context(0)
log_n, widths = 20, (84, 135, 20, 16)
cfg = pkg.FriConfig()
fp = cfg.fri_params(log_n)
gen = torch.Generator(device="cuda:0"); gen.manual_seed(1)
batches = []
for w in widths:
    d = torch.randint(0, 1 << 62, (w, 1 << log_n), dtype=torch.int64, device="cuda:0", generator=gen)
    torch.cuda.synchronize()
    batches.append(pkg.PolynomialBatch.from_values(ctx, None, cfg.rate_bits, False, cfg.cap_height, device_ptr=d.data_ptr(), shape=(w, 1 << log_n)))
    del d
zeta = splitmix_field(3, (2,))
gz = np.array([int(zeta[0]) * 7 % pkg.P, int(zeta[1]) * 7 % pkg.P], dtype=np.uint64)
polys = [[(o, j) for o, w in enumerate(widths) for j in range(w)], [(2, 0), (2, 1)]]
inst = pkg.FriInstanceInfo(np.stack([zeta, gz]), polys)
for rep in range(3):
    ch = pkg.Challenger(ctx)
    if rep == 2:
        torch.cuda.cudart().cudaProfilerStart()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    st = torch.cuda.ExternalStream(ctx.stream, device=0)
    e0.record(st)
    proof = pkg.prove_openings(inst, batches, ch, fp)
    e1.record(st)
    torch.cuda.synchronize()
    if rep == 2:
        torch.cuda.cudart().cudaProfilerStop()
    print("prove_openings %.3f ms" % e0.elapsed_time(e1))
