#!/usr/bin/env python3
"""End-to-end commit time from PINNED host memory (bench.py's `e2e`) for several upload schedules: one fresh process per
BSX_CHUNK_SCHEDULE value given on the command line ("default" = the library's own)."""
import os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if len(sys.argv) > 1 and sys.argv[1] == "--child":
    sys.path.insert(0, ROOT)
    import numpy as np, torch
    import __graft_entry__ as G
    from bench import splitmix_field, LOG_N, N_COLS
    pkg = G.load_package()
    ctx = pkg.Context(0)
    stream = torch.cuda.ExternalStream(ctx.stream, device=0)
    n = 1 << LOG_N
    host = torch.empty((N_COLS, n), dtype=torch.int64).pin_memory()
    a = host.numpy().view(np.uint64)
    a[:] = splitmix_field(0x5EED0001, (N_COLS, n))
    ts = []
    for rep in range(9):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        b = pkg.PolynomialBatch.from_values(ctx, a, 3, False, 4)
        e1.record(stream)
        torch.cuda.synchronize()
        cap0 = int(b.cap[0, 0])
        b.close()
        ts.append(e0.elapsed_time(e1))
    print("%-40s e2e %.2f ms (median of 6)  cap0=%016x" % (os.environ.get("BSX_CHUNK_SCHEDULE", "default"), float(np.median(ts[3:])), cap0))
else:
    for sched in sys.argv[1:] or ["default"]:
        env = dict(os.environ)
        env.pop("BSX_CHUNK_SCHEDULE", None)
        if sched != "default":
            env["BSX_CHUNK_SCHEDULE"] = sched
        r = subprocess.run([sys.executable, os.path.abspath(__file__), "--child"], env=env, capture_output=True, text=True)
        print(r.stdout.strip() or "FAILED: " + r.stderr[-300:], flush=True)
