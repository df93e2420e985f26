#!/usr/bin/env python3
"""Times bsx_batch_export of the row-major leaves (what a CPU-side quotient evaluation would download)."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import __graft_entry__ as G
pkg = G.load_package()
ctx = pkg.Context(0)
log_n, cols, rb = int(sys.argv[1]) if len(sys.argv) > 1 else 20, 135, 3
d = torch.randint(0, 1 << 62, (cols, 1 << log_n), dtype=torch.int64, device="cuda:0")
torch.cuda.synchronize()
b = pkg.PolynomialBatch.from_values(ctx, None, rb, False, 4, device_ptr=d.data_ptr(), shape=(cols, 1 << log_n))
nb = 1 << (log_n + rb)
gb = 8.0 * nb * cols / 1e9
pageable = np.empty((nb, cols), dtype=np.uint64)
pageable[:] = 0  # touch the pages
pinned_t = torch.empty((nb, cols), dtype=torch.int64).pin_memory()
pinned = pinned_t.numpy().view(np.uint64)
for name, buf in (("pageable", pageable), ("pinned", pinned)):
    for rep in range(3):
        t0 = time.perf_counter()
        ctx.check(pkg.lib().bsx_batch_export(b._h, None, buf.ctypes.data, None))
        dt = time.perf_counter() - t0
    print("export leaves 2^%d x %d (%.2f GB) to %s host memory: %.1f ms, %.1f GB/s" % (log_n + rb, cols, gb, name, dt * 1e3, gb / dt))
assert (pageable == pinned).all()
