"""Several prover threads in one process (plonky2x map-reduce runs independent proofs concurrently): one bsx_ctx per host
thread on the same GPU, no shared mutable state between contexts.  ctypes releases the GIL during the calls, so the
library really runs concurrently: staged uploads through the process-wide copy pool, per-context twiddle caches and
buffer caches, the allocation-free entry point's shared scratch behind the caller's own serialisation."""
import threading

import numpy as np
import pytest

from conftest import splitmix_field
import __graft_entry__ as G

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def pkg():
    return G.load_package()


def _work(pkg, ctx, seed):
    """One thread's job: two commits (pageable vectors -> staged upload; contiguous array) and a FRI proof."""
    log_n, n_cols = 13, 24
    vals = splitmix_field(seed, (n_cols, 1 << log_n))
    vecs = [vals[c].copy() for c in range(n_cols)]
    a = pkg.PolynomialBatch.from_values(ctx, vecs, 2, False, 3)
    b = pkg.PolynomialBatch.from_values(ctx, vals[:5], 2, False, 3)
    cfg = pkg.FriConfig(rate_bits=2, cap_height=3, proof_of_work_bits=6, num_query_rounds=5)
    fp = cfg.fri_params(log_n)
    zeta = splitmix_field(seed + 1, (1, 2))
    inst = pkg.FriInstanceInfo(zeta, [[(0, j) for j in range(n_cols)] + [(1, j) for j in range(5)]])
    ch = pkg.Challenger(ctx)
    ch.observe_cap(a.cap)
    ch.observe_cap(b.cap)
    proof = pkg.prove_openings(inst, [a, b], ch, fp)
    out = (a.cap.copy(), b.cap.copy(), a.get_lde_values(3, 1).copy(), proof.copy())
    a.close()
    b.close()
    return out


def test_four_contexts_on_four_threads(pkg):
    n_threads, reps = 4, 3
    ref_ctx = pkg.Context(0)
    want = [_work(pkg, ref_ctx, 7000 + 10 * t) for t in range(n_threads)]
    ref_ctx.close()
    errors, results = [], [None] * n_threads

    def run(t):
        try:
            ctx = pkg.Context(0)
            for _ in range(reps):
                results[t] = _work(pkg, ctx, 7000 + 10 * t)
                for got, exp in zip(results[t], want[t]):
                    assert (got == exp).all(), "thread %d: result differs from the single-threaded run" % t
            ctx.close()
        except Exception as ex:  # surfaced in the main thread
            errors.append("thread %d: %r" % (t, ex))

    threads = [threading.Thread(target=run, args=(t,)) for t in range(n_threads)]
    for th in threads:
        th.start()
    for th in threads:
        th.join(timeout=300)
    assert not any(th.is_alive() for th in threads), "a prover thread hung"
    assert not errors, errors
