#!/usr/bin/env python3
"""Device time of the quotient evaluation (bsx_quotient_values) at the headline shape: constants+sigmas (84 / 86 columns),
wires (135) and Zs + partial products (20) committed from random columns at 2^20 rows, rate_bits 3; the kernel reads
the three resident LDEs (15.3 GB) and writes 2 x 2^23 quotient values."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import __graft_entry__ as G
from bench import prove_circuit, LOG_N
pkg = G.load_package()
ctx = pkg.Context(0)
stream = torch.cuda.ExternalStream(ctx.stream, device=0)
n = 1 << LOG_N
gen = torch.Generator(device="cuda:0"); gen.manual_seed(5)
def commit(w):
    t = torch.randint(0, 1 << 62, (w, n), dtype=torch.int64, device="cuda:0", generator=gen)
    torch.cuda.synchronize()
    b = pkg.PolynomialBatch.from_values(ctx, None, 3, False, 4, device_ptr=t.data_ptr(), shape=(w, n))
    del t
    return b
wires, zs = commit(135), commit(20)
out = torch.empty((2, n << 3), dtype=torch.int64, device="cuda:0")
bytes_ = 8 * (84 + 135 + 20 + 2) * (n << 3) + 16 * (n << 3)
for gate_set in ("basic", "all"):
    circuit = prove_circuit(pkg, gate_set)
    cs = commit(circuit.num_constants + 80)
    ts = []
    for rep in range(6):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        pkg.quotient_values(ctx, circuit, cs, wires, zs, [1, 2, 3, 4], [3, 4], [5, 6], [7, 8], out_device_ptr=out.data_ptr())
        e1.record(stream)
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ms = float(np.median(ts[1:]))
    qb = [pkg.compute_quotient_polys(ctx, circuit, cs, wires, zs, [1, 2, 3, 4], [3, 4], [5, 6], [7, 8], 4) for _ in range(4)]
    print("gates %-5s (%d kinds, %d constraints): kernels alone (events inside bsx_compute_quotient_polys) %s ms" %
          (gate_set, len(circuit.gates), circuit.num_gate_constraints, [round(b.eval_ms, 2) for b in qb]))
    print("   quotient_values %.2f ms  (%.0f GB/s of LDE rows read; checksum %016x)" %
          (ms, bytes_ / ms / 1e6, int(out.view(-1)[12345].item()) & (2**64 - 1)))
    for b in qb:
        b.close()
    cs.close()
