#!/usr/bin/env python3
"""bench.py -- LDE + Poseidon-Merkle commit throughput (BASELINE.json `metric`) on N B200s.

A step is one `PolynomialBatch::from_values` over one synthetic witness: iNTT, coset LDE, leaf
hashing, Merkle tree down to the cap.  Workload at every N: the headline configuration
2^20 rows x 135 columns, rate_bits 3, cap_height 4 per GPU (one independent commit per rank: the
header_range map jobs of the north star share no state, so ranks exchange nothing on the data path
-- weak scaling).  `value` = committed input elements (N * C per rank, all ranks) per second with
the witness already in HBM; `e2e` = the same through the C ABI with the witness in pinned HOST
memory (H2D inside the timed call, cap back on the host).

At N > 1 the line's `value` / `e2e` are ONE commit of the same shape split over the N GPUs (`"scaling": "strong"`:
column-sharded transforms whose last pass stores into the owners' row slabs over NVLink P2P, a device-side flag
barrier, per-GPU leaf hashing and cap subtrees, caps exchanged through mailboxes behind the slabs; see
0-kno-blobstreamx_b200/csrc/api_shard.cu bsx_shard_run_p2p).  The N independent commits (weak scaling) move to
`replicas`, and `sharded` holds every exchange mode (p2p, p2p with a host barrier, staged NCCL all-to-all, p2p from
pinned host memory).

`prove` (skip with --no-prove): wall seconds of one shape-equivalent synthetic proof per GPU (three commits
from pinned host memory, the opening set, prove_openings) -- BASELINE.json's second metric on random columns.

`roofline` is the leaf-hash kernel (72 % of the step) against the measured HBM copy rate, as the contract asks; its
`issue` member is the bound that actually applies to Poseidon: warp-instruction issue slots (static SASS counts from
profiles/*_sass_cost.json x the measured permutation rate, against 148 SMs x 4 schedulers x SM clock).

`--impl reference` times the CPU arm instead: the in-repo oracle port of the plonky2 algorithm
(the reference mount holds no source and there is no Rust toolchain) with all
host threads on a bounded sample of the same workload.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")  # stdout carries exactly one JSON line
P = 0xFFFFFFFF00000001

LOG_N, N_COLS, RATE_BITS, CAP_HEIGHT = 20, 135, 3, 4
# --impl reference: a bounded sample per step (same columns / rate / cap, 2^18 rows: about 10 s of host time per
# step on 16 cores, so K = 20 steps end within a few minutes).  Its `config` names the sample, not the GPU workload.
REF_SAMPLE_LOG_N = int(os.environ.get("BSX_REF_SAMPLE_LOG_N", "18"))


def splitmix_field(seed, shape, offset=0):
    """Counter-based SplitMix64 stream reduced mod p; `offset` starts the counter there, so a slice of a
    larger array can be generated on its own (the column ranges of the sharded commit)."""
    n = int(np.prod(shape))
    out = np.empty(n, dtype=np.uint64)
    chunk = 1 << 24
    with np.errstate(over="ignore"):
        for lo in range(0, n, chunk):
            idx = np.arange(offset + lo + 1, offset + min(lo + chunk, n) + 1, dtype=np.uint64)
            z = np.uint64(seed) + idx * np.uint64(0x9E3779B97F4A7C15)
            z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
            z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
            z = z ^ (z >> np.uint64(31))
            out[lo:lo + idx.size] = z % np.uint64(P)
    return out.reshape(shape)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, device):
        self.device, self.rows, self.proc = device, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.device), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
                for nm, v in zip(names, r[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(nm)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def measured_peak_gbs():
    try:
        d = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def measured_traffic():
    """DRAM bytes per launch from the newest committed `ncu --set full` capture (profiles/*_traffic.json)."""
    import glob
    import re

    def version(path):  # r01_v10_traffic.json after r01_v8_traffic.json
        m = re.search(r"r(\d+)_v(\d+)", os.path.basename(path))
        return (int(m.group(1)), int(m.group(2))) if m else (0, 0)
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "*_traffic.json")), key=version)
    if not files:
        return {}, None
    try:
        return json.load(open(files[-1])), os.path.basename(files[-1])
    except Exception:
        return {}, None


def sass_cost():
    """Static instructions per permutation of the shipped leaf-hash kernel (tools/sass_cost.py, profiles/*_sass_cost.json)."""
    import glob
    import re

    def version(path):
        m = re.search(r"r(\d+)_v(\d+)", os.path.basename(path))
        return (int(m.group(1)), int(m.group(2))) if m else (0, 0)
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "*_sass_cost.json")), key=version)
    try:
        return json.load(open(files[-1]))["per_permutation"], os.path.basename(files[-1])
    except Exception:
        return None, None


def issue_roofline(n_perms, ms, sm_mhz):
    """The bound that applies to Poseidon: warp-instruction issue slots.  One slot per instruction, an IMAD.WIDE
    worth three; peak = 148 SMs x 4 schedulers x SM clock."""
    cost, src = sass_cost()
    if not cost:
        return None
    clk = (sm_mhz or 1965.0) * 1e6
    peak = 148 * 4 * clk
    warps = n_perms / 32.0 / (ms * 1e-3)
    inst, slots = warps * cost["instructions"], warps * cost["issue_slots_n_plus_2w"]
    return {"bound": "issue", "unit": "G warp-instruction slots/s", "peak": peak / 1e9, "achieved": slots / 1e9,
            "frac": slots / peak, "instructions_issued_frac": inst / peak,
            "instructions_per_permutation": cost["instructions"], "wide_multiplies_per_permutation": cost["wide"],
            "source": src, "sm_mhz": clk / 1e6,
            "note": "static SASS counts x measured permutation rate; instructions_issued_frac is what ncu reports "
                    "as smsp__issue_active (70.4 % in profiles/r02_v5_leafhash_ncu_summary.txt, sm__throughput 83.8 %)"}


def ntt_issue_roofline(traffic, n_in, n_out, ifft_ms, lde_ms, sm_mhz):
    """The bound that applies to the transforms, same method as the Poseidon one: warp instructions issued (per-element
    counts from the committed ncu capture x the elements of the live launch) / live time, against 148 SMs x 4 schedulers x
    SM clock; `slots` weighs an IMAD.WIDE as three."""
    ins = (traffic or {}).get("ntt_instructions")
    if not ins:
        return None
    clk = (sm_mhz or 1965.0) * 1e6
    peak = 148 * 4 * clk
    w = 1.0 + 2.0 * ins.get("wide_multiply_share", 0.0)

    def one(per_elem, n, ms):
        warps = per_elem * n / 32.0 / (ms * 1e-3)
        return {"instructions_per_element": per_elem, "instructions_issued_frac": warps / peak, "frac": warps * w / peak}
    return {"bound": "issue", "unit": "fraction of 592 schedulers x SM clock", "sm_mhz": clk / 1e6,
            "ifft": one(ins["ifft_pass_hi_per_element"] + ins["ifft_pass_lo_per_element"], n_in, ifft_ms),
            "lde": one(ins["lde_pass_hi_per_output_element"] + ins["lde_pass_lo_per_output_element"], n_out, lde_ms),
            "source": ins.get("source")}


def cpu_reference_arm(args, rank, world):
    """--impl reference: the oracle port on all host threads, one bounded sample of the workload per step.  The
    line's `config` describes the sample that was timed (2^REF_SAMPLE_LOG_N rows), not the 2^20-row GPU workload."""
    if rank != 0:
        return
    from oracle import oracle as O
    n = 1 << REF_SAMPLE_LOG_N
    vals = splitmix_field(0x5EED0001, (N_COLS, n))
    # torchrun exports OMP_NUM_THREADS=1 to every worker; the reference arm is meant to use the whole host
    O.set_num_threads(len(os.sched_getaffinity(0)))
    cores = O.num_threads()
    warm = min(args.warmup, 1)  # a CPU commit has no clocks or caches to warm beyond the first touch of its buffers
    for _ in range(warm):
        O.batch_from_values(vals, RATE_BITS, CAP_HEIGHT, want=())
    t0 = time.perf_counter()
    for _ in range(args.steps):
        O.batch_from_values(vals, RATE_BITS, CAP_HEIGHT, want=())
    dt = (time.perf_counter() - t0) / args.steps
    v = n * N_COLS / dt / 1e6
    sample = "2^%d rows x %d cols per step (1/%d of the 2^%d-row GPU workload), rate_bits %d, cap_height %d" % (
        REF_SAMPLE_LOG_N, N_COLS, 1 << (LOG_N - REF_SAMPLE_LOG_N), LOG_N, RATE_BITS, CAP_HEIGHT)
    cfg = workload_config(args.gpus, REF_SAMPLE_LOG_N)
    cfg["sharding"] = "rank 0 only, all host threads"
    cfg["l2"] = "host run"
    print(json.dumps({
        "impl": "reference", "metric": "lde_poseidon_merkle_commit", "value": v, "unit": "Melem/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": warm, "ms_per_step": dt * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
        "config": cfg,
        "cpu_baseline": {"value": v, "unit": "Melem/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": "Melem/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "note": "scalar OpenMP C restatement of the plonky2 algorithm (oracle/), NOT the plonky2 crate: the reference "
                "mount holds no source and no Rust toolchain exists on this image.  plonky2's AVX2 / AVX-512 Poseidon "
                "would move this number up severalfold, so a ratio against it is not a speed-up over the reference "
                "prover.  Throughput in Melem/s is nearly size independent on the CPU (2^15: 3.6, 2^18: see value), "
                "so the sample stands for the 2^20 workload; bench.py's own arm times the full 2^20 oracle commit "
                "once as `cpu_baseline`.",
    }))


def workload_config(n_gpus, log_n=LOG_N):
    what = "per GPU" if log_n == LOG_N else "per step: a 1/%d sample of the 2^%d-row headline workload" % (
        1 << (LOG_N - log_n), LOG_N)
    return {"workload": "PolynomialBatch::from_values commit, 2^%d rows x %d cols %s, rate_bits %d, "
                        "cap_height %d (BASELINE.json configs[4] headline point; standard_recursion_config)"
                        % (log_n, N_COLS, what, RATE_BITS, CAP_HEIGHT),
            "log_n": log_n, "n_cols": N_COLS, "rate_bits": RATE_BITS, "cap_height": CAP_HEIGHT,
            "sharding": "one independent commit per GPU, no data-path collective" if n_gpus > 1 else "single GPU",
            "l2": "inputs (1.13 GB) and LDE (9.06 GB) exceed the 126 MB L2; no flush needed"}


def sharded_arm(args, pkg, ctx, stream, rank, world, local_rank, cap_rank0):
    """Strong-scaling leg at N > 1 -- the line's `value` there: ONE 2^20 x 135 commit sharded over the N GPUs (columns
    -> ranks for the transforms, leaf rows -> ranks for hashing).  Exchange modes: "p2p" = the last NTT pass stores into
    the owners' row slabs over NVLink and the ranks meet at a device-side flag barrier (no NCCL call, one host sync per
    commit); "p2p_host" = the same stores with a host barrier (torch.distributed) and an NCCL cap gather; "nccl" =
    staged all-to-all.  "p2p" is also timed end to end: every rank uploads ITS columns from pinned host memory inside
    the timed call.  The witness is rank 0's witness of the independent-commit leg, so the cap must equal rank 0's cap."""
    import torch
    import torch.distributed as dist
    n = 1 << LOG_N
    dev = "cuda:%d" % local_rank
    plan = pkg.ShardPlan(LOG_N, N_COLS, RATE_BITS, CAP_HEIGHT, world)
    c0, c1 = plan.column_range(rank)
    host = torch.empty((c1 - c0, n), dtype=torch.int64).pin_memory()
    host_np = host.numpy().view(np.uint64)
    host_np[:] = splitmix_field(0x5EED0001, (c1 - c0, n), offset=c0 * n)
    loc = host.to(dev)
    want = torch.from_numpy(np.ascontiguousarray(cap_rank0).view(np.int64).copy()).to(dev)
    dist.broadcast(want, 0)
    want = want.cpu().numpy().view(np.uint64)
    out = {"scaling": "strong", "workload": "one 2^%d x %d commit over %d GPUs" % (LOG_N, N_COLS, world)}
    with torch.cuda.stream(stream):
        for mode in ("p2p", "p2p_host", "nccl", "p2p_e2e"):
            sb = pkg.ShardedPolynomialBatch.setup(ctx, LOG_N, N_COLS, RATE_BITS, CAP_HEIGHT, mode=mode.split("_e2e")[0])

            def step(timing=False):
                if mode == "p2p_e2e":
                    sb.run(host_np, timing=timing)
                else:
                    sb.run(None, device_ptr=loc.data_ptr(), timing=timing)
            for _ in range(max(2, min(args.warmup, 3))):
                step()
            assert (sb.cap == want).all(), "sharded cap (%s) differs from the single-GPU cap" % mode
            dist.barrier()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            tr = []
            for _ in range(args.steps):
                step(True)
                tr.append(sb.timing.as_dict())
            e1.record(stream)
            dist.barrier()
            torch.cuda.synchronize()
            assert (sb.cap == want).all(), "sharded cap (%s) differs from the single-GPU cap" % mode
            keys = ["ifft_ms", "lde_ms", "h2d_ms", "leaf_hash_ms", "tree_ms"]
            t = torch.tensor([e0.elapsed_time(e1) / args.steps] + [float(np.mean([x[k] for x in tr])) for k in keys],
                             device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t[0])
            out[mode] = {"ms_per_commit": ms, "value": n * N_COLS / (ms * 1e-3) / 1e6, "unit": "Melem/s",
                         "transforms_ms_max": float(t[1]) + float(t[2]),
                         "exchange_bytes_per_gpu": 8 * (c1 - c0) * (world - 1) * plan.chunk}
            if mode.startswith("p2p") and mode != "p2p_host":
                out[mode].update({"peer_wait_ms_max": float(t[3]), "leaf_hash_ms_max": float(t[4]), "tree_and_cap_exchange_ms_max": float(t[5])})
            if mode == "p2p_e2e":
                out[mode]["h2d_bytes_per_step_all_ranks"] = 8 * n * N_COLS
            sb.close()
    del loc
    return out


PROVE_WIDTHS = (86, 135, 20, 16)  # constants (4 selectors + 2) + sigmas | wires | Zs+partial products | quotient chunks (8 x 2 challenges)


def prove_circuit(pkg, gate_set="all"):
    """standard_recursion_config shape for the synthetic prove: 135 wires, 80 routed, and the gate table of a
    recursion-style circuit in four selector groups (what gates/selectors.rs forms for max degree 9): [Noop, Constant,
    PublicInput, PoseidonMds, BaseSum<2>, Reducing], [ReducingExtension, Arithmetic, ArithmeticExtension, MulExtension,
    Exponentiation], [RandomAccess(4 bits), CosetInterpolation(4 bits, degree 6)], [Poseidon]; "basic": the four base-field gates only.  The committed columns
    are random, so the gate and permutation constraints do not vanish on H; with quotient_degree_factor = 8 = 2^3
    `trim_to_len` checks nothing (upstream's prover would not notice either), and the work per proof is exactly that of a
    satisfied circuit."""
    g = pkg.Gate
    if gate_set == "basic":
        gates = [g(pkg.GATE_NOOP, 0, 0, 0, 0, 2), g(pkg.GATE_CONSTANT, 2, 1, 0, 0, 2),
                 g(pkg.GATE_PUBLIC_INPUT, 0, 2, 1, 2, 4), g(pkg.GATE_ARITHMETIC, 20, 3, 1, 2, 4)]
        return pkg.CircuitDescription(LOG_N, gates, num_constants=4, num_selectors=2)
    gates = [g(pkg.GATE_NOOP, 0, 0, 0, 0, 6), g(pkg.GATE_CONSTANT, 2, 1, 0, 0, 6), g(pkg.GATE_PUBLIC_INPUT, 0, 2, 0, 0, 6),
             g(pkg.GATE_POSEIDON_MDS, 0, 3, 0, 0, 6), g(pkg.GATE_BASE_SUM, 32, 4, 0, 0, 6), g(pkg.GATE_REDUCING, 43, 5, 0, 0, 6),
             g(pkg.GATE_REDUCING_EXTENSION, 32, 6, 1, 6, 11), g(pkg.GATE_ARITHMETIC, 20, 7, 1, 6, 11),
             g(pkg.GATE_ARITHMETIC_EXTENSION, 10, 8, 1, 6, 11), g(pkg.GATE_MUL_EXTENSION, 13, 9, 1, 6, 11),
             g(pkg.GATE_EXPONENTIATION, 66, 10, 1, 6, 11), g(pkg.GATE_RANDOM_ACCESS, 4 | 4 << 8 | 2 << 16, 11, 2, 11, 13),
             g(pkg.GATE_COSET_INTERPOLATION, 4 | 6 << 8, 12, 2, 11, 13), g(pkg.GATE_POSEIDON, 0, 13, 3, 13, 14)]
    return pkg.CircuitDescription(LOG_N, gates, num_constants=6, num_selectors=4)


def prove_arm(args, pkg, ctx, stream, rank, world, local_rank, wires_np):
    """BASELINE.json's second metric, as far as it can exist here: the commitment-phase call sequence of ONE plonky2
    proof at 2^20 rows, standard_recursion_config, on random columns (the plonky2x circuits and their witness
    generators are not in the mount: SURVEY.md 8(d)).  Per proof: commit wires (from_values, pinned host), commit
    Zs (from_values), commit quotient chunks (from_coeffs), evaluate the opening set at zeta / g*zeta, prove_openings
    (alpha-combine, FRI commit rounds, proof of work, 28 query rounds) -> proof words on host.  The constants+sigmas
    batch is circuit data and is committed before the clock starts.  Every rank runs its own proofs (header_range
    map jobs: replicas, no communication); the time is device time on the library's stream, max over ranks."""
    import torch
    import torch.distributed as dist
    n = 1 << LOG_N
    cfg = pkg.FriConfig()
    fp = cfg.fri_params(LOG_N)

    def pinned(seed, w):
        t = torch.empty((w, n), dtype=torch.int64).pin_memory()
        a = t.numpy().view(np.uint64)
        a[:] = splitmix_field(seed + rank, (w, n))
        return t, a

    keep0, cs = pinned(0x5EED0200, PROVE_WIDTHS[0])
    keep2, zs = pinned(0x5EED0300, PROVE_WIDTHS[2])
    constants_sigmas = pkg.PolynomialBatch.from_values(ctx, cs, cfg.rate_bits, False, cfg.cap_height)
    g = pow(7, (P - 1) >> LOG_N, P)
    circuit = prove_circuit(pkg)
    pis = splitmix_field(0x5EED0600 + rank, (4,))
    eval_ms = []

    debug = os.environ.get("BSX_BENCH_DEBUG") == "1"
    nbytes = [0]

    def one_proof():
        marks = [("start", time.perf_counter())]

        def mark(name):
            if debug:
                ctx.sync()
                marks.append((name, time.perf_counter()))

        ch = pkg.Challenger(ctx)
        ch.observe_cap(constants_sigmas.cap)
        wires = pkg.PolynomialBatch.from_values(ctx, wires_np, cfg.rate_bits, False, cfg.cap_height)
        mark("commit_wires")
        ch.observe_cap(wires.cap)
        bg = ch.get_n_challenges(4)  # betas, gammas
        zb = pkg.PolynomialBatch.from_values(ctx, zs, cfg.rate_bits, False, cfg.cap_height)
        mark("commit_zs")
        ch.observe_cap(zb.cap)
        alphas = ch.get_n_challenges(2)
        # compute_quotient_polys on the device from the three resident LDEs, coset_ifft, chunks, commitment
        qb = pkg.compute_quotient_polys(ctx, circuit, constants_sigmas, wires, zb, pis, bg[:2], bg[2:], alphas,
                                        cfg.cap_height)
        eval_ms.append(qb.eval_ms)
        mark("quotient")
        ch.observe_cap(qb.cap)
        zeta = ch.get_extension_challenge()
        gzeta = np.array([int(zeta[0]) * g % P, int(zeta[1]) * g % P], dtype=np.uint64)
        batches = [constants_sigmas, wires, zb, qb]
        openings = [b.eval_commitment(zeta) for b in batches] + [zb.eval_commitment(gzeta, 0, 2)]
        mark("opening_set")
        for o in openings:
            ch.observe_elements(o.reshape(-1))
        polys0 = [(o, j) for o, w in enumerate(PROVE_WIDTHS) for j in range(w)]
        inst = pkg.FriInstanceInfo(np.array([zeta, gzeta]), [polys0, [(2, 0), (2, 1)]])
        proof = pkg.prove_openings(inst, batches, ch, fp)
        mark("prove_openings")
        # ProofWithPublicInputs::to_bytes (the opening set in OpeningSet field order; 4 constant columns)
        cs_o, w_o, z_o, q_o, zn_o = openings
        pw = pkg.ProofWithPublicInputs(pkg.Proof(wires.cap, zb.cap, qb.cap, pkg.OpeningSet(
            cs_o[:4], cs_o[4:], w_o, z_o[:2], zn_o, z_o[2:], q_o), proof), pis)
        nbytes[0] = len(pw.to_bytes(fp, list(PROVE_WIDTHS)))
        mark("proof_bytes")
        for b in batches[1:]:
            b.close()
        mark("free")
        if debug:
            print("rank %d proof phases ms: %s" % (rank, ", ".join(
                "%s %.1f" % (marks[i][0], (marks[i][1] - marks[i - 1][1]) * 1e3) for i in range(1, len(marks)))),
                file=sys.stderr)
        return proof

    first = one_proof()
    one_proof()  # second warm-up: the stream-ordered pool is still growing during the first two proofs
    reps = max(3, min(args.steps, 5))
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    times = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        proof = one_proof()
        e1.record(stream)
        torch.cuda.synchronize()
        times.append(e0.elapsed_time(e1) / 1e3)
    assert (proof == first).all(), "identical proofs differ"
    s_per = float(np.median(times))
    s_max = max(times)
    if world > 1:
        t = torch.tensor([s_per, s_max], device="cuda:%d" % local_rank, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        s_per, s_max = float(t[0]), float(t[1])
    # the two gate-independent prover steps either side of quotient evaluation, on device-resident inputs (not part
    # of wall_s_per_proof: the synthetic proof above commits random Zs / quotient chunks from host memory)
    device_steps = None
    try:
        n_routed, n_chunks = 80, 10
        gen = torch.Generator(device="cuda:%d" % local_rank)
        gen.manual_seed(0x5EED0500 + rank)
        dw = torch.randint(0, 1 << 62, (n_routed, n), dtype=torch.int64, device="cuda:%d" % local_rank, generator=gen)
        ds = torch.randint(0, 1 << 62, (n_routed, n), dtype=torch.int64, device="cuda:%d" % local_rank, generator=gen)
        dz = torch.empty((2 * n_chunks, n), dtype=torch.int64, device="cuda:%d" % local_rank)
        dq = torch.randint(0, 1 << 62, (2, n << 3), dtype=torch.int64, device="cuda:%d" % local_rank, generator=gen)
        k_is = np.array([pow(7, j, P) for j in range(n_routed)], dtype=np.uint64)
        tz, tq = [], []
        for _ in range(4):
            e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
            e[0].record(stream)
            ctx.partial_products_and_zs(None, None, k_is, [3, 4], [5, 6], 8, out_device_ptr=dz.data_ptr(),
                                        device_ptrs=(dw.data_ptr(), ds.data_ptr()), shape=(n_routed, n))
            e[1].record(stream)
            qb = pkg.PolynomialBatch.from_quotient_values(ctx, None, LOG_N, 8, cfg.rate_bits, False, cfg.cap_height,
                                                          device_ptr=dq.data_ptr(), shape=(2, n << 3))
            e[2].record(stream)
            torch.cuda.synchronize()
            qb.close()
            tz.append(e[0].elapsed_time(e[1]))
            tq.append(e[1].elapsed_time(e[2]))
        device_steps = {"partial_products_and_zs_ms": float(np.median(tz[1:])),
                        "from_quotient_values_ms": float(np.median(tq[1:])),
                        "workload": "80 routed wires x 2^%d rows, 2 challenges, max_degree 8 -> 20 columns; 2 x 2^%d quotient "
                                    "values -> coset_ifft -> 16 chunks -> commitment; inputs resident in HBM" % (LOG_N, LOG_N + 3)}
        del dw, ds, dz, dq
    except Exception as ex:  # an auxiliary measurement must not take the headline down
        device_steps = {"error": str(ex)[:200]}
    constants_sigmas.close()
    del keep0, keep2
    return {"wall_s_per_proof": s_per, "device_steps": device_steps, "wall_s_per_proof_max": s_max, "proofs_per_s": world / s_per, "n_gpus": world,
            "reps": reps, "statistic": "median of reps, max over ranks",
            "proof_words": int(proof.size), "proof_with_public_inputs_bytes": nbytes[0],
            "quotient_eval_ms": float(np.median(eval_ms[2:])) if len(eval_ms) > 2 else None,
            "fri_arity_bits": list(fp.reduction_arity_bits),
            "workload": "shape-equivalent synthetic prove, 2^%d rows: commits of %d/%d columns from host memory, the "
                        "quotient evaluated on the device from the resident LDEs (14 gate kinds incl. PoseidonGate's 123 constraints, 4 selector groups) and "
                        "committed as %d chunks (+%d columns precommitted), opening set at zeta and g*zeta, "
                        "prove_openings, ProofWithPublicInputs bytes; standard_recursion_config; random columns, "
                        "NOT a real header_range proof; one proof per GPU, replicas" %
                        (LOG_N, PROVE_WIDTHS[1], PROVE_WIDTHS[2], PROVE_WIDTHS[3], PROVE_WIDTHS[0])}


def bind_to_gpu_numa_node(local_rank):
    """What `numactl --cpunodebind` does in a deployment with one rank per GPU: run this process (and the library's copy
    threads, which inherit the mask) on the CPUs NVML lists as local to its GPU, so that first-touch places the pinned and
    pageable host buffers on that socket's memory.  Returns what was done for the JSON line; any failure leaves the
    process unbound."""
    try:
        import pynvml
        import torch
        pynvml.nvmlInit()
        pr = torch.cuda.get_device_properties(local_rank)
        bus = "%08x:%02x:%02x.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)
        try:
            h = pynvml.nvmlDeviceGetHandleByPciBusId(bus)
        except TypeError:
            h = pynvml.nvmlDeviceGetHandleByPciBusId(bus.encode())
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
        cpus = [64 * w + b for w, v in enumerate(words) for b in range(64) if (int(v) >> b) & 1]
        allowed = sorted(set(cpus) & set(os.sched_getaffinity(0)))
        if not allowed:
            return {"bound": False, "why": "no local CPU in the allowed set"}
        os.sched_setaffinity(0, allowed)
        return {"bound": True, "gpu": bus, "cpus": len(allowed), "first_cpu": allowed[0]}
    except Exception as e:  # no NVML, no permission, an old torch: measured unbound
        return {"bound": False, "why": "%s: %s" % (type(e).__name__, str(e)[:80])}


def gpu_arm(args, rank, world, local_rank):
    # stdout carries exactly one JSON line: whatever libraries print there (NCCL's version banner) goes to stderr
    real_stdout = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    import torch
    import torch.distributed as dist
    import __graft_entry__ as G
    pkg = G.load_package()
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    torch.cuda.set_device(local_rank)
    # N > 1 only: at N = 1 the CPU baseline leg of this process wants every host core
    numa = (bind_to_gpu_numa_node(local_rank) if world > 1 and os.environ.get("BSX_BENCH_NO_NUMA_BIND") != "1"
            else {"bound": False, "why": "single rank" if world == 1 else "disabled"})
    ctx = pkg.Context(local_rank)  # raises without the built CUDA library or a GPU: no fallback
    stream = torch.cuda.ExternalStream(ctx.stream, device=local_rank)
    n = 1 << LOG_N
    host = torch.empty((N_COLS, n), dtype=torch.int64).pin_memory()
    host_np = host.numpy().view(np.uint64)
    host_np[:] = splitmix_field(0x5EED0001 + rank, (N_COLS, n))
    dev = host.to("cuda:%d" % local_rank)
    torch.cuda.synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step_device(timing):
        b = pkg.PolynomialBatch.from_values(ctx, None, RATE_BITS, False, CAP_HEIGHT, device_ptr=dev.data_ptr(),
                                            shape=(N_COLS, n), timing=timing)
        t = b.timing.as_dict() if timing else None
        cap = b.cap
        b.close()
        return cap, t

    def step_host():
        b = pkg.PolynomialBatch.from_values(ctx, host_np, RATE_BITS, False, CAP_HEIGHT)
        cap = b.cap
        b.close()
        return cap

    cap0 = None
    for _ in range(args.warmup):
        cap0, _ = step_device(True)
    parity_rows = parity_paths = None
    if world == 1 and not args.no_cpu_baseline:  # openings of the commit the oracle re-does on the host below
        b = pkg.PolynomialBatch.from_values(ctx, None, RATE_BITS, False, CAP_HEIGHT, device_ptr=dev.data_ptr(), shape=(N_COLS, n))
        probe = [0, 0x2AAAAA, (n << RATE_BITS) - 1]
        parity_rows = b.merkle_tree.get(np.array(probe, dtype=np.uint64))
        parity_paths = [b.merkle_tree.prove(i) for i in probe]
        b.close()
    # ---- device-resident arm
    launches0 = pkg.lib().bsx_launch_count()
    sampler = ClockSampler(local_rank)
    barrier()
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    phases = []
    for _ in range(args.steps):
        cap, t = step_device(True)
        phases.append(t)
    e1.record(stream)
    barrier()
    clocks = sampler.stop()
    ms = e0.elapsed_time(e1) / args.steps
    launches = (pkg.lib().bsx_launch_count() - launches0) // args.steps
    assert (cap == cap0).all(), "cap changed between identical steps"
    # ---- end-to-end arm: pinned host witness -> cap on host, through the C ABI
    for _ in range(max(1, min(args.warmup, 2))):
        cap_h = step_host()
    assert (cap_h == cap0).all(), "host-buffer path and device-buffer path disagree"
    barrier()
    h0, h1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    h0.record(stream)
    for _ in range(args.steps):
        step_host()
    h1.record(stream)
    barrier()
    ms_e2e = h0.elapsed_time(h1) / args.steps
    # ---- the same from PAGEABLE memory, one separately allocated vector per column (what a Rust
    # Vec<PolynomialValues<F>> is): host threads stage the columns through the context's pinned ring
    vecs = [host_np[c].copy() for c in range(N_COLS)]
    cap_p = None
    for _ in range(2):
        b = pkg.PolynomialBatch.from_values(ctx, vecs, RATE_BITS, False, CAP_HEIGHT)
        cap_p = b.cap
        b.close()
    assert (cap_p == cap0).all(), "pageable column-vector path disagrees"
    barrier()
    p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    p0.record(stream)
    n_p = max(1, min(args.steps, 3))
    for _ in range(n_p):
        pkg.PolynomialBatch.from_values(ctx, vecs, RATE_BITS, False, CAP_HEIGHT).close()
    p1.record(stream)
    barrier()
    ms_pageable = p0.elapsed_time(p1) / n_p
    del vecs
    if world > 1:
        t = torch.tensor([ms, ms_e2e, ms_pageable], device="cuda:%d" % local_rank, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, ms_e2e, ms_pageable = float(t[0]), float(t[1]), float(t[2])
    sharded = sharded_arm(args, pkg, ctx, stream, rank, world, local_rank, cap0) if world > 1 else None
    del dev
    torch.cuda.empty_cache()
    prove = None if args.no_prove else prove_arm(args, pkg, ctx, stream, rank, world, local_rank, host_np)
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    elems = n * N_COLS
    nb = n << RATE_BITS
    value = world * elems / (ms * 1e-3) / 1e6
    e2e = world * elems / (ms_e2e * 1e-3) / 1e6
    avg = {k: float(np.mean([p[k] for p in phases])) for k in phases[0]}
    peak, peak_src = measured_peak_gbs()

    def roof(bytes_, ms_, traffic=None):
        a = bytes_ / (ms_ * 1e-3) / 1e9
        return {"bound": "hbm", "achieved": a, "peak": peak, "unit": "GB/s", "frac": a / peak, "traffic": traffic,
                "frac_of_nominal_8000": a / 8000.0}  # against the measured copy peak and the nominal one

    traffic, traffic_src = measured_traffic()
    leaf_bytes = 8 * nb * N_COLS + 32 * nb
    roofline = roof(leaf_bytes, avg["leaf_hash_ms"], traffic.get("leaf_hash_kernel"))
    roofline["traffic_source"] = traffic_src
    # the bound that actually applies (integer pipes), from the same committed ncu capture
    roofline["pipes"] = (traffic.get("pipes") or {}).get("leaf_hash_kernel")
    roofline["issue"] = issue_roofline(nb * ((N_COLS + 7) // 8), avg["leaf_hash_ms"], (clocks or {}).get("sm_mhz"))
    roofline.update({"kernel": "leaf_hash_kernel", "ms_per_launch": avg["leaf_hash_ms"], "peak_source": peak_src,
                     "algorithmic_bytes": leaf_bytes,
                     "note": "dominant kernel by time; Poseidon is integer-pipe bound, so its HBM fraction is low "
                             "by construction (see DESIGN.md and profiles/ for pipe utilisation)"})
    perms = nb * ((N_COLS + 7) // 8) + nb - (1 << CAP_HEIGHT)
    out = {
        "metric": "lde_poseidon_merkle_commit", "value": value, "unit": "Melem/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u64", "data": "synthetic", "config": workload_config(world),
        "lde_melem_per_s": value * (1 << RATE_BITS),
        "e2e": {"value": e2e, "unit": "Melem/s", "ms_per_step": ms_e2e, "h2d_bytes_per_step": 8 * elems,
                "d2h_bytes_per_step": 32 << CAP_HEIGHT},
        "e2e_pageable": {"value": world * elems / (ms_pageable * 1e-3) / 1e6, "unit": "Melem/s", "ms_per_step": ms_pageable,
                         "source": "%d separately allocated pageable host vectors per commit (Vec<PolynomialValues<F>>), "
                                   "staged by host threads through a pinned ring" % N_COLS},
        "gpu_launches": int(launches), "clocks": clocks, "host_binding": numa, "roofline": roofline,
        "phases_ms": avg,
        "roofline_ntt": {
            "ifft": dict(roof(16 * elems, avg["ifft_ms"], traffic.get("ifft")), algorithmic_bytes=16 * elems),
            "lde": dict(roof(8 * elems + 8 * nb * N_COLS, avg["lde_ms"], traffic.get("lde")),
                        algorithmic_bytes=8 * elems + 8 * nb * N_COLS),
            "pipes": {k: v for k, v in (traffic.get("pipes") or {}).items() if k.startswith("ntt")},
            "issue": ntt_issue_roofline(traffic, elems, nb * N_COLS, avg["ifft_ms"], avg["lde_ms"], (clocks or {}).get("sm_mhz")),
            "note": "two kernels each (pass HI + pass LO); bound by integer issue / dispatch slots, not by HBM: a Goldilocks "
                    "butterfly is ~33 integer instructions per 16 bytes moved, ~500 instructions per element (DESIGN.md 3.4); "
                    "`issue` is the bound that applies, the HBM fractions are low by construction",
        },
        "whole_commit": dict(roof(8 * elems * (2 + (1 << RATE_BITS)) + 64 * nb, ms),
                             algorithmic_bytes=8 * elems * (2 + (1 << RATE_BITS)) + 64 * nb),
        "poseidon": {"permutations_per_step": perms,
                     "gperm_per_s": perms / ((avg["leaf_hash_ms"] + avg["tree_ms"]) * 1e-3) / 1e9},
    }
    if sharded is not None:
        # N > 1: the headline of the line is ONE commit sharded over the N GPUs (strong scaling, the north star's
        # "column-sharded NTT + NVLink exchange + per-GPU subtrees"); the N independent commits stay as `replicas`
        out["replicas"] = {"scaling": "weak", "value": out["value"], "ms_per_step": out["ms_per_step"], "e2e": out["e2e"],
                           "e2e_pageable": out["e2e_pageable"], "lde_melem_per_s": out["lde_melem_per_s"],
                           "workload": "one independent 2^%d x %d commit per GPU (header_range map jobs)" % (LOG_N, N_COLS)}
        sh, she = sharded["p2p"], sharded["p2p_e2e"]
        out.update({"value": sh["value"], "ms_per_step": sh["ms_per_commit"], "scaling": "strong",
                    "lde_melem_per_s": sh["value"] * (1 << RATE_BITS),
                    "e2e": {"value": she["value"], "unit": "Melem/s", "ms_per_step": she["ms_per_commit"],
                            "h2d_bytes_per_step": 8 * elems, "d2h_bytes_per_step": world * world * (32 << max(CAP_HEIGHT - (world.bit_length() - 1), 0))}})
        del out["e2e_pageable"]
        out["config"]["sharding"] = ("ONE commit over %d GPUs: columns -> ranks for iNTT + LDE, the last NTT pass stores into the "
                                     "owners' row slabs over NVLink P2P, device-side flag barrier, per-GPU leaf hashing and cap "
                                     "subtrees, caps exchanged through the slabs' mailboxes (no NCCL on the data path)" % world)
        out["config"]["workload"] = out["config"]["workload"].replace(" per GPU", ", one commit over all GPUs")
        out["sharded"] = sharded
    if prove is not None:
        out["prove"] = prove
    if world == 1 and not args.no_cpu_baseline:
        # The CPU arm at the TRUE size, once: it is both the timed baseline and the parity check of the commit the
        # line reports (same witness, same shape): cap, one LDE row and its Merkle path against the oracle.
        from oracle import oracle as O
        O.set_num_threads(len(os.sched_getaffinity(0)))
        t0 = time.perf_counter()
        want = O.batch_from_values(host_np, RATE_BITS, CAP_HEIGHT, want=("leaves", "digests"))
        dt = time.perf_counter() - t0
        log_nb = LOG_N + RATE_BITS
        probe = [0, 0x2AAAAA, nb - 1]
        rows_ok = all((parity_rows[k] == want["leaves"][i]).all() for k, i in enumerate(probe))
        paths_ok = all((parity_paths[k] == O.merkle_prove(want["digests"], log_nb, CAP_HEIGHT, i)).all()
                       for k, i in enumerate(probe))
        out["parity"] = {"config": "2^%d x %d, rate_bits %d, cap_height %d (the timed workload)" % (LOG_N, N_COLS, RATE_BITS, CAP_HEIGHT),
                         "against": "oracle/ (CPU restatement of plonky2; parity unpinned against the real prover)",
                         "cap_equal": bool((cap0 == want["cap"]).all()), "lde_rows_equal": bool(rows_ok),
                         "merkle_paths_equal": bool(paths_ok), "leaves_probed": probe}
        del want
        out["cpu_baseline"] = {
            "value": elems / dt / 1e6, "unit": "Melem/s", "cores": O.num_threads(), "kind": "port", "seconds": dt,
            "sample": "the whole 2^%d rows x %d cols commit, once (rate_bits %d, cap_height %d), scalar OpenMP C oracle; "
                      "a CPU restatement, not the plonky2 crate (no AVX2/AVX-512 Poseidon)" % (LOG_N, N_COLS, RATE_BITS, CAP_HEIGHT)}
        assert out["parity"]["cap_equal"] and rows_ok and paths_ok, "device commit differs from the oracle at the headline size"
    print(json.dumps(out), file=real_stdout, flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-prove", action="store_true", help="skip the shape-equivalent synthetic prove leg")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 0)
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        cpu_reference_arm(args, rank, world)
        return
    if args.warmup < 3:
        args.warmup = 3
    gpu_arm(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
