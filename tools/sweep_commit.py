#!/usr/bin/env python3
"""BASELINE.json configs[4]: LDE + Merkle commit sweep over rows x columns x rate_bits on one GPU.
Device-resident random witnesses (canonical field elements drawn on the device), CUDA-event timing of
`PolynomialBatch::from_values`, best of `--reps`.  Points whose buffers would not fit `--max-gb` are skipped
and listed.  Every measured point is CHECKED on its last commit (`"checked": true`): three opened leaf rows verify
against the cap with the oracle's `merkle_verify` (CPU), the committed polynomials evaluated at 7 * omega_NB^i equal
LDE row i (ties the leaves to the coefficients), and evaluated at omega_N^i they equal input row i (ties the
coefficients to the witness).  Output: one JSON object per point (profiles/r02_sweep.jsonl)."""
import argparse, json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import __graft_entry__ as G
ap = argparse.ArgumentParser()
ap.add_argument("--log-n", default="16,18,20,22,24")
ap.add_argument("--cols", default="80,135,256,400")
ap.add_argument("--rate-bits", default="1,2,3,4")
ap.add_argument("--reps", type=int, default=2)
ap.add_argument("--max-gb", type=float, default=165.0)
a = ap.parse_args()
pkg = G.load_package()
from oracle import oracle as O  # checker only: merkle_verify of opened rows on the CPU
P = 0xFFFFFFFF00000001
W32 = 1753635133440165772  # plonky2 POWER_OF_TWO_GENERATOR (order 2^32)


def root(log_n):
    return pow(W32, 1 << (32 - log_n), P)


def bitrev(x, bits):
    return int(format(x, "0%db" % bits)[::-1], 2) if bits else 0


def check_point(b, dev, log_n, rb, cols):
    """the three checks of the module docstring on batch `b` committed from the device tensor `dev`"""
    log_nb, nb = log_n + rb, 1 << (log_n + rb)
    cap = b.cap
    rng = np.random.default_rng(log_n * 1000 + cols * 10 + rb)
    for leaf in (0, int(rng.integers(0, nb)), nb - 1):
        row, path = b.merkle_tree.get(leaf), b.merkle_tree.prove(leaf)
        if not O.merkle_verify(row, leaf, path, cap):
            return False
        i = bitrev(leaf, log_nb)  # natural LDE index of that leaf
        x = 7 * pow(root(log_nb), i, P) % P
        ev = b.eval_commitment(np.array([x, 0], dtype=np.uint64))
        if not ((ev[:, 0] == row[:cols]).all() and (ev[:, 1] == 0).all()):
            return False
    for i in (0, int(rng.integers(0, 1 << log_n)), (1 << log_n) - 1):
        x = pow(root(log_n), i, P)
        ev = b.eval_commitment(np.array([x, 0], dtype=np.uint64))
        want = dev[:, i].cpu().numpy().view(np.uint64) % np.uint64(P)
        if not ((ev[:, 0] == want).all() and (ev[:, 1] == 0).all()):
            return False
    return True


ctx = pkg.Context(0)
stream = torch.cuda.ExternalStream(ctx.stream, device=0)
gen = torch.Generator(device="cuda:0")
gen.manual_seed(0x5EED0005)
for log_n in [int(x) for x in a.log_n.split(",")]:
    n = 1 << log_n
    for cols in [int(x) for x in a.cols.split(",")]:
        dev = None
        for rb in [int(x) for x in a.rate_bits.split(",")]:
            nb = n << rb
            need = 8 * (2 * n * cols + nb * cols) + 64 * nb + 8 * (n << rb)  # witness + coeffs + LDE + digests + table
            if need > a.max_gb * 1e9:
                print(json.dumps({"log_n": log_n, "cols": cols, "rate_bits": rb, "skipped": "needs %.0f GB" % (need / 1e9)}), flush=True)
                continue
            if dev is None:
                dev = torch.randint(0, 1 << 62, (cols, n), dtype=torch.int64, device="cuda:0", generator=gen)
                torch.cuda.synchronize()  # torch filled it on ITS stream; the library works on the context's stream
            best, cap0 = None, None
            for rep in range(a.reps + 1):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream)
                b = pkg.PolynomialBatch.from_values(ctx, None, rb, False, 4, device_ptr=dev.data_ptr(), shape=(cols, n), timing=True)
                e1.record(stream)
                torch.cuda.synchronize()
                t, ph = e0.elapsed_time(e1), b.timing.as_dict()
                c = int(b.cap[0, 0])
                assert cap0 in (None, c)
                cap0 = c
                if rep == a.reps:
                    checked = check_point(b, dev, log_n, rb, cols)
                b.close()
                if rep and (best is None or t < best[0]):
                    best = (t, ph)
            assert checked, "sweep point %d x %d r%d failed its correctness checks" % (log_n, cols, rb)
            print(json.dumps({"log_n": log_n, "cols": cols, "rate_bits": rb, "ms": round(best[0], 3),
                              "melem_per_s": round(n * cols / best[0] / 1e3, 1),
                              "lde_melem_per_s": round(nb * cols / best[0] / 1e3, 1),
                              "phases_ms": {k: round(v, 3) for k, v in best[1].items()}, "checked": bool(checked)}), flush=True)
        del dev
        torch.cuda.empty_cache()
