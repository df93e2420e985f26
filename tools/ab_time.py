#!/usr/bin/env python3
"""A/B timing of library builds: for every libbsx_*.so given, a fresh process (BSX_B200_LIB) commits the headline
shape `--reps` times from device memory and prints the median phase times and the first cap word (all builds must
agree).  tools/ab_time.py build/variants/libbsx_ntt_*.so"""
import argparse, json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ap = argparse.ArgumentParser()
ap.add_argument("libs", nargs="*")
ap.add_argument("--log-n", type=int, default=20)
ap.add_argument("--cols", type=int, default=135)
ap.add_argument("--rate-bits", type=int, default=3)
ap.add_argument("--reps", type=int, default=7)
ap.add_argument("--child", action="store_true")
a = ap.parse_args()
if a.child:
    sys.path.insert(0, ROOT)
    import numpy as np
    import torch
    import __graft_entry__ as G
    from bench import splitmix_field
    pkg = G.load_package()
    ctx = pkg.Context(0)
    n = 1 << a.log_n
    dev = torch.from_numpy(splitmix_field(0x5EED0001, (a.cols, n)).view(np.int64)).to("cuda:0")
    torch.cuda.synchronize()
    ph, cap0 = [], None
    for rep in range(a.reps + 2):
        b = pkg.PolynomialBatch.from_values(ctx, None, a.rate_bits, False, 4, device_ptr=dev.data_ptr(), shape=(a.cols, n), timing=True)
        cap0 = int(b.cap[0, 0])
        if rep >= 2:
            ph.append(b.timing.as_dict())
        b.close()
    med = {k: round(float(np.median([p[k] for p in ph])), 3) for k in ph[0]}
    print(json.dumps({"lib": os.path.basename(pkg.library_path()), "cap0": "%016x" % cap0, **med}))
else:
    for lib in a.libs or [""]:
        env = dict(os.environ)
        if lib:
            env["BSX_B200_LIB"] = os.path.abspath(lib)
        r = subprocess.run([sys.executable, os.path.abspath(__file__), "--child", "--log-n", str(a.log_n), "--cols", str(a.cols),
                            "--rate-bits", str(a.rate_bits), "--reps", str(a.reps)], env=env, capture_output=True, text=True)
        print(r.stdout.strip() or ("FAILED %s: %s" % (lib, r.stderr[-400:])), flush=True)
