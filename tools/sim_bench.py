#!/usr/bin/env python
"""epv_simulate (forward simulation on the device) at the headline shard size, wall clock of
the whole C-ABI call pair (simulate + fetch to host arrays), beside the unmodified reference
binaries (oracle/_ref/bin/epievo_sim + global_jumps_to_paths, one core) on a sample of the same
tree.  Prints one JSON line."""
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "oracle")]

import ctypes as C  # noqa: E402

import numpy as np  # noqa: E402

import cpu_bench  # noqa: E402
from epievo_b200.device import Context  # noqa: E402

wl = sys.argv[1] if len(sys.argv) > 1 else "bal8"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 12_500_000
tree = cpu_bench.make_tree(wl)
rates, T = cpu_bench.test_model()
ctx = Context(0)
ctx.set_tree(tree); ctx.set_model(rates, T)
ctx.simulate(100_000, 1, tree.branches)                     # warm-up
nj = C.c_uint64()
t0 = time.time()
ctx._ck(ctx.lib.epv_simulate(ctx.h, n, None, 2, C.byref(nj)))
t1 = time.time()
launches = ctx.launches() if hasattr(ctx, "launches") else None
ps = None
init = np.zeros(n * tree.n_nodes, np.uint8); off = np.zeros(n * tree.n_nodes + 1, np.uint64)
jumps = np.zeros(max(1, nj.value))
t2 = time.time()
ctx._ck(ctx.lib.epv_simulate_fetch(ctx.h, init.ctypes.data_as(C.POINTER(C.c_uint8)),
                                   off.ctypes.data_as(C.POINTER(C.c_uint64)),
                                   jumps.ctypes.data_as(C.POINTER(C.c_double))))
t3 = time.time()
n_br = tree.n_nodes - 1
out = {"workload": wl, "n_sites": n, "n_branches": n_br, "n_jumps": int(nj.value),
       "jumps_per_row": nj.value / (n * n_br),
       "gpu": {"simulate_s": t1 - t0, "fetch_s": t3 - t2, "rows_per_s": n * n_br / (t1 - t0)}}
ctx.close()
BIN = os.path.join(ROOT, "oracle", "_ref", "bin")
GOLD = os.path.join(ROOT, "tests", "golden", "cli")
if os.path.exists(BIN + "/epievo_sim") and wl == "testtree":
    m = 200_000
    with tempfile.TemporaryDirectory() as tmp:
        t0 = time.time()
        subprocess.run([BIN + "/epievo_sim", "-s", "1", "-n", str(m), "-t", GOLD + "/tree.nwk", "-p", tmp + "/j",
                        GOLD + "/test.param", tmp + "/s"], check=True, capture_output=True)
        t1 = time.time()
        subprocess.run([BIN + "/global_jumps_to_paths", "-t", GOLD + "/tree.nwk", tmp + "/s", tmp + "/j", tmp + "/p"],
                       check=True, capture_output=True)
        t2 = time.time()
    out["reference_1core"] = {"sample_sites": m, "epievo_sim_s": t1 - t0, "global_jumps_to_paths_s": t2 - t1,
                              "rows_per_s": m * n_br / (t2 - t0),
                              "note": "binaries incl. their text output; the Gillespie loop draws one event of the whole sequence at a time"}
print(json.dumps(out))
