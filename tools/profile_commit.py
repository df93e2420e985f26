#!/usr/bin/env python3
"""One or more commits at a chosen shape, for ncu captures (kept short: ncu replays every kernel)."""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as G  # noqa: E402
from bench import splitmix_field  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--log-n", type=int, default=20)
ap.add_argument("--cols", type=int, default=135)
ap.add_argument("--rate-bits", type=int, default=3)
ap.add_argument("--cap-height", type=int, default=4)
ap.add_argument("--reps", type=int, default=2)
a = ap.parse_args()
pkg = G.load_package()
ctx = pkg.Context(0)
vals = splitmix_field(0x5EED0001, (a.cols, 1 << a.log_n))
for i in range(a.reps):
    b = pkg.PolynomialBatch.from_values(ctx, vals, a.rate_bits, False, a.cap_height, timing=True)
    print(i, b.timing.as_dict(), "cap0=%016x" % int(b.cap[0, 0]))
    b.close()
ctx.close()
