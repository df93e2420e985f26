"""Worker of test_device_barrier_refuses_two_processes_on_one_gpu: rank r of 2, both on cuda:0, IPC handles exchanged through
files in a directory given on the command line.  Prints the status of bsx_shard_run_p2p and the library's message."""
import ctypes as C
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as G  # noqa: E402

rank, d = int(sys.argv[1]), sys.argv[2]
pkg = G.load_package()
ctx = pkg.Context(0)
plan = pkg.ShardPlan(10, 6, 1, 1, 2)
sb = pkg.ShardedPolynomialBatch.create(ctx, plan, rank)
open(os.path.join(d, "h%d.tmp" % rank), "wb").write(sb.ipc_handle())
os.rename(os.path.join(d, "h%d.tmp" % rank), os.path.join(d, "h%d" % rank))
peer = os.path.join(d, "h%d" % (1 - rank))
t0 = time.time()
while not os.path.exists(peer):
    if time.time() - t0 > 60:
        sys.exit("peer handle never arrived")
    time.sleep(0.05)
sb.set_peer_ipc(1 - rank, open(peer, "rb").read())
c0, c1 = plan.column_range(rank)
vals = np.arange((c1 - c0) << 10, dtype=np.uint64).reshape(c1 - c0, 1 << 10)
caps = np.zeros((2, 1, 4), dtype=np.uint64)
h = C.c_void_p()
f = pkg.lib().bsx_shard_run_p2p
f.restype, f.argtypes = C.c_int32, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.POINTER(C.c_void_p), C.c_void_p, C.c_void_p]
rc = f(sb._h, vals.ctypes.data, 0, 0, C.byref(h), caps.ctypes.data, None)
print("rc=%d msg=%s" % (rc, pkg.lib().bsx_last_error(ctx._h).decode()), flush=True)
# keep the slab mapped until the peer has made its call too
open(os.path.join(d, "done%d" % rank), "w").write("x")
t0 = time.time()
while not os.path.exists(os.path.join(d, "done%d" % (1 - rank))) and time.time() - t0 < 60:
    time.sleep(0.05)
