#!/usr/bin/env python3
"""Shape-equivalent synthetic prove at standard_recursion_config: the four PolynomialBatch commits of one plonky2
proof (constants+sigmas, wires, Z/partial products, quotient chunks) followed by prove_openings (alpha-combine,
divide by (X - z), FRI commit rounds, proof of work, 28 query rounds).  Random columns stand in for the witness:
BASELINE.json configs 2-4 need the plonky2x gadget stack, which does not exist here."""
import argparse, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as G
from bench import splitmix_field
ap = argparse.ArgumentParser()
ap.add_argument("--log-n", type=int, default=20)
ap.add_argument("--widths", default="84,135,20,16")
ap.add_argument("--reps", type=int, default=3)
a = ap.parse_args()
pkg = G.load_package()
ctx = pkg.Context(0)
n = 1 << a.log_n
widths = [int(w) for w in a.widths.split(",")]
cfg = pkg.FriConfig()
fp = cfg.fri_params(a.log_n)
t0 = time.perf_counter()
hosts = [splitmix_field(0x5EED0100 + i, (w, n)) for i, w in enumerate(widths)]
t_gen = time.perf_counter() - t0
res = []
for rep in range(a.reps):
    t0 = time.perf_counter()
    batches = [pkg.PolynomialBatch.from_values(ctx, h, cfg.rate_bits, False, cfg.cap_height) for h in hosts]
    t_commit = time.perf_counter() - t0
    ch = pkg.Challenger(ctx)
    for b in batches:
        ch.observe_cap(b.cap)
    zeta = ch.get_extension_challenge()
    gzeta = np.array([(int(zeta[0]) * 7) % pkg.P, (int(zeta[1]) * 7) % pkg.P], dtype=np.uint64)
    polys0 = [(o, j) for o, w in enumerate(widths) for j in range(w)]
    polys1 = [(2, j) for j in range(min(2, widths[2]))] if len(widths) > 2 else []
    inst = pkg.FriInstanceInfo(np.array([zeta, gzeta]) if polys1 else np.array([zeta]), [polys0, polys1] if polys1 else [polys0])
    t0 = time.perf_counter()
    proof = pkg.prove_openings(inst, batches, ch, fp)
    t_open = time.perf_counter() - t0
    res.append((t_commit, t_open, int(proof[:4].sum() & 0xFFFF)))
    for b in batches:
        b.close()
print(json.dumps({"workload": "4 commits (%s cols) + prove_openings at 2^%d, standard_recursion_config, host witness" % (a.widths, a.log_n),
                  "commit_s": [round(r[0], 4) for r in res], "prove_openings_s": [round(r[1], 4) for r in res],
                  "proof_words": int(proof.size), "arity_bits": fp.reduction_arity_bits, "host_gen_s": round(t_gen, 2)}))
