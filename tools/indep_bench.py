#!/usr/bin/env python
"""epievo_initialization's site-independent E-step at the headline shard size: the three
O(n_sites) calls timed on the GPU (CUDA events, paths resident) beside the unmodified
reference (oracle/_ref, one core) on a sample.  Prints one JSON line."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "oracle")]

import numpy as np  # noqa: E402
import torch  # noqa: E402

import cpu_bench  # noqa: E402
import cpu_libs  # noqa: E402
from epievo_b200.device import Context  # noqa: E402
from epievo_b200.synthetic import simulate_paths  # noqa: E402

wl = sys.argv[1] if len(sys.argv) > 1 else "bal8"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 12_500_000
tree = cpu_bench.make_tree(wl)
rates, T = cpu_bench.test_model()
r2 = (1.3, 0.7)
ps = simulate_paths(tree, n, rates, T, seed=1)
ctx = Context(0)
st = torch.cuda.Stream(); torch.cuda.set_stream(st); ctx.set_stream(st.cuda_stream)
ctx.set_tree(tree); ctx.set_model(rates, T); ctx.upload_paths_compact(ps)


def timed(f, reps):
    f()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st)
    for _ in range(reps):
        f()
    e1.record(st)
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


it = [0]


def upd():
    ctx.indep_update_paths(r2, 5, it[0]); it[0] += 1


out = {"workload": wl, "n_sites": n, "n_branches": tree.n_nodes - 1,
       "gpu_ms": {"expectation": timed(lambda: ctx.indep_expectation(r2), 5),
                  "suffstats": timed(ctx.indep_suffstats, 5),
                  "update_paths": timed(upd, 3)}}
ctx.close()
if cpu_libs.ref_available():
    m = 50_000
    sub = ps.site_slice(0, m)
    ref = cpu_libs.ref()
    ref.indep_load(sub)
    cpu = {}
    t0 = time.perf_counter(); ref.indep_expectation(tree, r2); cpu["expectation"] = time.perf_counter() - t0
    t0 = time.perf_counter(); ref.indep_suffstats(); cpu["suffstats"] = time.perf_counter() - t0
    t0 = time.perf_counter(); ref.indep_update_paths(tree, r2, 1); cpu["update_paths"] = time.perf_counter() - t0
    out["cpu_reference_s_per_%d_sites" % m] = cpu
    out["speedup_vs_1_core"] = {k: cpu[k] * (n / m) / (out["gpu_ms"][k] * 1e-3) for k in cpu}
print(json.dumps(out))
