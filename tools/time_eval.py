#!/usr/bin/env python3
"""Times bsx_batch_eval_ext (opening-set evaluation) on a committed batch: wall time per call through the C ABI
(includes the 16-byte H2D, the launches and the D2H of the results)."""
import argparse
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as G  # noqa: E402
from bench import splitmix_field  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--log-n", type=int, default=20)
ap.add_argument("--cols", type=int, default=135)
ap.add_argument("--rate-bits", type=int, default=1)
ap.add_argument("--reps", type=int, default=10)
a = ap.parse_args()
pkg = G.load_package()
ctx = pkg.Context(0)
vals = splitmix_field(0x5EED0001, (a.cols, 1 << a.log_n))
b = pkg.PolynomialBatch.from_values(ctx, vals, a.rate_bits, False, 4)
for k in (1, 2):
    pts = splitmix_field(5 + k, (k, 2))
    b.eval_commitment(pts)
    t0 = time.perf_counter()
    for _ in range(a.reps):
        b.eval_commitment(pts)
    dt = (time.perf_counter() - t0) / a.reps
    gb = 8.0 * a.cols * (1 << a.log_n) / 1e9
    print("eval_commitment 2^%d x %d, %d point(s): %.3f ms per call, %.0f GB/s of coefficients" %
          (a.log_n, a.cols, k, dt * 1e3, gb / dt))
b.close()
ctx.close()
