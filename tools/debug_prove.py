#!/usr/bin/env python3
"""Per-phase timing of the synthetic prove loop with the memory pool's footprint (debugging aid)."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import __graft_entry__ as G
from bench import splitmix_field, prove_circuit, LOG_N, N_COLS, P
pkg = G.load_package()
ctx = pkg.Context(0)
n = 1 << LOG_N
mode = sys.argv[1] if len(sys.argv) > 1 else "device"
def pinned(seed, w):
    t = torch.empty((w, n), dtype=torch.int64).pin_memory()
    a = t.numpy().view(np.uint64); a[:] = splitmix_field(seed, (w, n)); return t, a
k0, cs = pinned(1, 84); k1, wires_np = pinned(2, 135); k2, zs = pinned(3, 20); k3, qs = pinned(4, 16)
cfg = pkg.FriConfig()
constants_sigmas = pkg.PolynomialBatch.from_values(ctx, cs, 3, False, 4)
circuit = prove_circuit(pkg)
pis = splitmix_field(9, (4,))
def free_gb():
    return torch.cuda.mem_get_info(0)[0] / 1e9
for it in range(8):
    t = [time.perf_counter()]
    f0 = free_gb()
    wires = pkg.PolynomialBatch.from_values(ctx, wires_np, 3, False, 4, timing=True); ctx.sync(); t.append(time.perf_counter())
    tw = wires.timing.as_dict()
    zb = pkg.PolynomialBatch.from_values(ctx, zs, 3, False, 4); ctx.sync(); t.append(time.perf_counter())
    if mode == "device":
        qb = pkg.compute_quotient_polys(ctx, circuit, constants_sigmas, wires, zb, pis, [3, 4], [5, 6], [7, 8], 4)
    else:
        qb = pkg.PolynomialBatch.from_coeffs(ctx, qs, 3, False, 4)
    ctx.sync(); t.append(time.perf_counter())
    f1 = free_gb()
    for b in (wires, zb, qb):
        b.close()
    ctx.sync(); t.append(time.perf_counter())
    print(mode, it, "ms:", [round((t[i + 1] - t[i]) * 1e3, 1) for i in range(len(t) - 1)], "free GB before/after %.1f %.1f" % (f0, f1),
          "used %.2f" % (ctx.mem_used() / 1e9), {k: round(v, 1) for k, v in tw.items()}, flush=True)
