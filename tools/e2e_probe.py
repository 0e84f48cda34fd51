#!/usr/bin/env python
"""Where the end-to-end leg of bench.py spends its time: epv_upload_paths_compact, epv_run_mcmc and
epv_download_paths_compact timed one by one (host clock around a synchronised call), pinned host buffers.
usage: python tools/e2e_probe.py [--workload pairwise] [--sites N] [--iters 5]"""
import argparse
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import bench                                       # noqa: E402
from epievo_b200.device import Context             # noqa: E402
from epievo_b200.synthetic import simulate_paths   # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="pairwise")
    ap.add_argument("--sites", type=int, default=0)
    ap.add_argument("--iters", type=int, default=5)
    a = ap.parse_args()
    n = a.sites or bench.DEFAULT_SITES[a.workload]
    tree = bench.make_tree(a.workload)
    rates, T = bench.test_model()
    ps = simulate_paths(tree, n, rates, T, seed=12345)
    n_rows = n * tree.n_nodes
    jcap = int(ps.jumps.size * 1.5) + 65536
    h_jmp = torch.empty(jcap, dtype=torch.float64, pin_memory=True)
    h_code = torch.empty(n_rows, dtype=torch.int16, pin_memory=True)
    ctx = Context(0)
    st = torch.cuda.Stream()
    torch.cuda.set_stream(st)
    ctx.set_stream(st.cuda_stream)
    ctx.set_tree(tree); ctx.set_model(rates, T)
    ctx.upload_paths(ps); ctx.reset()
    tot = ps.tot_time.copy()
    ctx.run_mcmc(10, 10, seed=1, first_sweep=0)
    nj = ctx.download_compact_raw(h_code.data_ptr(), h_jmp.data_ptr(), jcap)
    rows = []
    sw = 20
    for _ in range(a.iters):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        ctx.upload_compact_raw(n, tree.n_nodes, h_code.data_ptr(), h_jmp.data_ptr(), nj, tot)
        torch.cuda.synchronize(); t1 = time.perf_counter()
        ctx.run_mcmc(10, 10, seed=1, first_sweep=sw); sw += 20
        torch.cuda.synchronize(); t2 = time.perf_counter()
        nj = ctx.download_compact_raw(h_code.data_ptr(), h_jmp.data_ptr(), jcap)
        torch.cuda.synchronize(); t3 = time.perf_counter()
        rows.append(((t1 - t0) * 1e3, (t2 - t1) * 1e3, (t3 - t2) * 1e3))
    r = np.array(rows)
    print("%s n=%d rows=%d jumps=%d recoveries=%d" % (a.workload, n, n_rows, nj, ctx.recovery_count()))
    print("ms per call (median of %d): upload %.3f  run_mcmc %.3f  download %.3f" % ((a.iters,) + tuple(np.median(r, 0))))
    print("all:", np.round(r, 3).tolist())


if __name__ == "__main__":
    main()
