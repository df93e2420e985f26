#!/usr/bin/env python3
"""Per-phase SM cycles of NTT pass HI (debug build: make -C 0-kno-blobstreamx_b200/csrc clean all EXTRA=-DNTT_PROFILE)."""
import ctypes as C
import sys, os
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as G
from bench import splitmix_field
pkg = G.load_package()
L = pkg.lib()
ctx = pkg.Context(0)
cols = int(sys.argv[1]) if len(sys.argv) > 1 else 135
log_n = int(sys.argv[2]) if len(sys.argv) > 2 else 20
rate_bits = int(sys.argv[3]) if len(sys.argv) > 3 else 3
vals = splitmix_field(1, (cols, 1 << log_n))
out = (C.c_ulonglong * 4)()
for rep in range(2):
    L.bsx_debug_ntt_profile(out, 1)
    b = pkg.PolynomialBatch.from_values(ctx, vals, rate_bits, False, 4, timing=True)
    L.bsx_debug_ntt_profile(out, 0)
    t = b.timing.as_dict()
    b.close()
tot = sum(out[:3])
print("2^%d x %d, rate_bits %d:" % (log_n, cols, rate_bits), end=" ")
print("pass HI cycles summed over CTAs (iNTT + LDE): load %.1f%%  stages %.1f%%  store %.1f%%   (lde %.2f ms, ifft %.2f ms)"
      % (100 * out[0] / tot, 100 * out[1] / tot, 100 * out[2] / tot, t["lde_ms"], t["ifft_ms"]))
