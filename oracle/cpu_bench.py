"""TEST / BENCH INFRASTRUCTURE ONLY -- CPU timing of the reference path.

Runs the reference's own SingleSiteSampler::reset + run_mcmc
(oracle/_ref/libepievo_ref.so = the unmodified reference sources, see oracle/Makefile;
falls back to the C restatement oracle/liboracle.so when that build is absent) on a
bounded sample of a benchmark workload, in-process, no file I/O in the timed region.
Imported only by bench.py (cpu_baseline leg and --impl reference).  The reference is
single-threaded; "all host cores" means P independent processes over disjoint site
blocks, which is how a user would parallelise it today.
"""
from __future__ import annotations

import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for _p in (ROOT, os.path.join(ROOT, "tests")):
    if _p not in sys.path:
        sys.path.insert(0, _p)


def make_tree(workload: str):
    from epievo_b200.paths import Tree, parse_newick
    if workload == "bal8":
        return Tree.balanced(8, 0.05)
    if workload == "bal32":
        return Tree.balanced(32, 0.05)
    if workload == "pairwise":
        return Tree.pair(1.0)
    if workload == "testtree":
        return parse_newick("((C:0.03,D:0.06)E:0.02,F:0.1)G:0.0;")
    raise ValueError("unknown workload " + workload)


def test_model():
    """reference test/test.param, rates scaled to one change per site per unit time."""
    import cpu_libs
    rates, T, _, _ = cpu_libs.oracle().model_from_params(0.844912, 0.893359, -0.8, -1.8)
    return rates, T


def worker(rank, workload, n_sites, burn_in, batch, n_steps, seed, barrier, out_q):
    """One process: its own site block, n_steps E-steps; reports per-step seconds."""
    import cpu_libs
    from epievo_b200.synthetic import simulate_paths
    cpu = cpu_libs.best_cpu()
    tree = make_tree(workload)
    rates, T = test_model()
    ps = simulate_paths(tree, n_sites, rates, T, seed + rank)
    cpu.load_paths(ps)
    cpu.seed(seed + rank)
    if barrier is not None:
        barrier.wait()
    times = []
    for _ in range(n_steps):
        t0 = time.perf_counter()
        cpu.run_mcmc(tree, rates, T, burn_in, batch, reset=True)
        times.append(time.perf_counter() - t0)
        if barrier is not None:
            barrier.wait()
    out_q.put((rank, cpu.kind, times))


def run(workload, n_sites, burn_in, batch, n_steps, procs, seed=777):
    """Returns (kind, per-step wall seconds as the max over processes)."""
    import multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    barrier = ctx.Barrier(procs) if procs > 1 else None
    ps = [ctx.Process(target=worker, args=(r, workload, n_sites, burn_in, batch, n_steps,
                                           seed, barrier, q)) for r in range(procs)]
    for p in ps:
        p.start()
    res = [q.get() for _ in ps]
    for p in ps:
        p.join()
    kind = res[0][1]
    per_step = [max(r[2][k] for r in res) for k in range(n_steps)]
    return kind, per_step
