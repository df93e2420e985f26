#!/usr/bin/env python3
"""Time `PolynomialBatch.from_quotient_values` (coset_ifft + chunks + from_coeffs) on device-resident quotient values:
n_polys x 2^(log_n + q) words, standard_recursion_config shape by default (2 challenges, factor 8, rate_bits 3)."""
import argparse, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import __graft_entry__ as G
ap = argparse.ArgumentParser()
ap.add_argument("--log-n", type=int, default=20)
ap.add_argument("--q-bits", type=int, default=3)
ap.add_argument("--factor", type=int, default=8)
ap.add_argument("--polys", type=int, default=2)
ap.add_argument("--reps", type=int, default=5)
a = ap.parse_args()
pkg = G.load_package()
ctx = pkg.Context(0)
stream = torch.cuda.ExternalStream(ctx.stream, device=0)
big = 1 << (a.log_n + a.q_bits)
gen = torch.Generator(device="cuda:0"); gen.manual_seed(9)
vals = torch.randint(0, 1 << 62, (a.polys, big), dtype=torch.int64, device="cuda:0", generator=gen)
torch.cuda.synchronize()
ts, tc = [], []
for rep in range(a.reps + 2):
    e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
    e0.record(stream)
    b = pkg.PolynomialBatch.from_quotient_values(ctx, None, a.log_n, a.factor, 3, False, 4, device_ptr=vals.data_ptr(),
                                                 shape=(a.polys, big))
    e1.record(stream)
    c = pkg.PolynomialBatch.from_coeffs(ctx, None, 3, False, 4, device_ptr=vals.data_ptr(),
                                        shape=(a.polys * a.factor, 1 << a.log_n))
    e2.record(stream)
    torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1)); tc.append(e1.elapsed_time(e2))
    b.close(); c.close()
print("from_quotient_values %d x 2^%d (factor %d): %.2f ms (best %.2f); from_coeffs of the same %d x 2^%d alone: %.2f ms"
      % (a.polys, a.log_n + a.q_bits, a.factor, float(np.median(ts[2:])), min(ts[2:]), a.polys * a.factor, a.log_n,
         float(np.median(tc[2:]))))
