#!/usr/bin/env python
"""Regenerates tests/golden/ from the UNMODIFIED reference build (oracle/_ref, made by
oracle/Makefile from /root/reference).  Run in the build container only; the outputs are
committed so that machines without /root/reference (the GPU box) can still pin the
oracle and the CUDA path against the reference's own results.

  cli/   the reference's README walk-through at n=400 with fixed seeds, run with the
         reference's own binaries: epievo_sim -> global_jumps_to_paths ->
         epievo_est_complete / epievo_est_params_histories (with and without -b)
  lib_*.npz   inputs + outputs of the reference library entry points (through
         oracle/ref_driver.cpp) on seeded synthetic histories
"""
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
for p in (ROOT, os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)

import cpu_libs                                             # noqa: E402
from epievo_b200.paths import Tree                          # noqa: E402
from epievo_b200.synthetic import simulate_paths            # noqa: E402
from synth import TEST_PARAMS, ref_test_tree, synth_paths   # noqa: E402

BIN = os.path.join(ROOT, "oracle", "_ref", "bin")
REF_TEST = "/root/reference/test"


def sh(cmd, **kw):
    return subprocess.run(cmd, check=True, capture_output=True, text=True, **kw)


def make_cli():
    out = os.path.join(HERE, "cli")
    os.makedirs(out, exist_ok=True)
    for f in ("tree.nwk", "test.param"):
        # tiny input fixtures of the reference (test/tree.nwk, test/test.param): values only
        open(os.path.join(out, f), "w").write(open(os.path.join(REF_TEST, f)).read())
    tree, param = os.path.join(out, "tree.nwk"), os.path.join(out, "test.param")
    sh([BIN + "/epievo_sim", "-s", "1", "-n", "400", "-t", tree, "-p", out + "/sim.jumps", param,
        out + "/sim.states"])
    sh([BIN + "/global_jumps_to_paths", "-t", tree, out + "/sim.states", out + "/sim.jumps",
        out + "/sim.paths"])
    sh([BIN + "/epievo_est_complete", "-o", out + "/est_complete.param", param, tree, out + "/sim.paths"])
    sh([BIN + "/epievo_est_complete", "-b", "-o", out + "/est_complete_b.param", "-t",
        out + "/est_complete_b.nwk", param, tree, out + "/sim.paths"])
    r = sh([BIN + "/epievo_est_params_histories", "-v", "-s", "1", "-i", "2", "-B", "3", "-L", "2",
            "-o", out + "/mcem.paths", "-p", out + "/mcem.param", param, tree, out + "/sim.paths"])
    open(out + "/mcem.log", "w").write("".join(l + "\n" for l in r.stderr.splitlines()
                                               if l[:1].isdigit() or l.startswith("itr")))
    r = sh([BIN + "/epievo_est_params_histories", "-v", "-b", "-s", "1", "-i", "2", "-B", "3", "-L", "2",
            "-o", out + "/mcem_b.paths", "-p", out + "/mcem_b.param", "-t", out + "/mcem_b.nwk",
            param, tree, out + "/sim.paths"])
    open(out + "/mcem_b.log", "w").write("".join(l + "\n" for l in r.stderr.splitlines()
                                                 if l[:1].isdigit() or l.startswith("itr")))
    # MCEM parameter trajectories of the CPU reference over seeds (for the drop-in CLI test)
    rows = []
    for seed in range(1, 13):
        sh([BIN + "/epievo_est_params_histories", "-s", str(seed), "-i", "3", "-B", "10", "-L", "10",
            "-o", out + "/_tmp.paths", "-p", out + "/_tmp.param", param, tree, out + "/sim.paths"])
        vals = [l.split()[1:] for l in open(out + "/_tmp.param")]
        rows.append("%d\t%s\t%s\t%s\t%s\n" % (seed, vals[0][0], vals[0][1], vals[1][0], vals[1][1]))
    open(out + "/mcem_seeds.tsv", "w").write("seed\tT00\tT11\tbaseline00\tbaseline11\n" + "".join(rows))
    os.remove(out + "/_tmp.paths")
    os.remove(out + "/_tmp.param")
    os.remove(out + "/sim.jumps")
    os.remove(out + "/mcem_b.paths")


def make_cli_b_seeds():
    """-b (branch lengths optimised too): final parameters AND branch lengths of the CPU reference binary
    over 12 seeds, `-b -i 3 -B 10 -L 10` on the walk-through data (for tests/test_gpu_mcem.py)."""
    from epievo_b200.paths import read_tree
    out = os.path.join(HERE, "cli")
    tree, param = os.path.join(out, "tree.nwk"), os.path.join(out, "test.param")
    rows = []
    for seed in range(1, 13):
        sh([BIN + "/epievo_est_params_histories", "-b", "-s", str(seed), "-i", "3", "-B", "10", "-L", "10",
            "-o", out + "/_tmp.paths", "-p", out + "/_tmp.param", "-t", out + "/_tmp.nwk", param, tree,
            out + "/sim.paths"])
        vals = [l.split()[1:] for l in open(out + "/_tmp.param")]
        br = read_tree(out + "/_tmp.nwk").branches[1:]
        rows.append("%d\t%s\t%s\t%s\t%s\t%s\n" % (seed, vals[0][0], vals[0][1], vals[1][0], vals[1][1],
                                                   "\t".join(repr(float(b)) for b in br)))
    open(out + "/mcem_b_seeds.tsv", "w").write("seed\tT00\tT11\tbaseline00\tbaseline11\tbranch1\tbranch2\tbranch3\tbranch4\n"
                                               + "".join(rows))
    for f in ("_tmp.paths", "_tmp.param", "_tmp.nwk"):
        os.remove(os.path.join(out, f))


def make_init():
    """epievo_initialization (site-independent EM + resampled starting paths) by the reference
    binary: the deterministic EM trajectory of -s 1, and the final parameters over 12 seeds."""
    out = os.path.join(HERE, "cli")
    tree, states = os.path.join(out, "tree.nwk"), os.path.join(out, "sim.states")
    r = sh([BIN + "/epievo_initialization", "-v", "-s", "1", "-i", "10", "-B", "5", "-o", out + "/_tmp.paths",
            "-p", out + "/_tmp.param", tree, states])
    em = [l for l in r.stderr.splitlines() if l.startswith("itr") or (l[:1].isdigit() and len(l.split()) == 3)]
    open(out + "/init.log", "w").write("".join(l + "\n" for l in em))
    rows = []
    for seed in range(1, 13):
        sh([BIN + "/epievo_initialization", "-s", str(seed), "-i", "10", "-B", "5", "-o", out + "/_tmp.paths",
            "-p", out + "/_tmp.param", tree, states])
        vals = [l.split()[1:] for l in open(out + "/_tmp.param")]
        rows.append("%d\t%s\t%s\t%s\t%s\n" % (seed, vals[0][0], vals[0][1], vals[1][0], vals[1][1]))
    open(out + "/init_seeds.tsv", "w").write("seed\tT00\tT11\tbaseline00\tbaseline11\n" + "".join(rows))
    os.remove(out + "/_tmp.paths")
    os.remove(out + "/_tmp.param")


CASES = {"testtree": (ref_test_tree, 300, 3.0), "pair": (lambda: Tree.pair(1.0), 300, 1.0),
         "bal8": (lambda: Tree.balanced(8, 0.05), 200, 2.0),
         "bal32": (lambda: Tree.balanced(32, 0.05), 60, 3.0)}
SITES = (1, 2, 29, 58)


def make_lib():
    ref = cpu_libs.ref()
    rates, T, _, _ = ref.model_from_params(*TEST_PARAMS)
    for name, (mk, n, dens) in CASES.items():
        tree = mk()
        ps = synth_paths(tree, n, 101, dens)
        ref.load_paths(ps)
        J, D = ref.get_suff_stats()
        tri = ref.tri_llh(rates, T)
        d = dict(n_sites=n, subtree_sizes=tree.subtree_sizes, parent_ids=tree.parent_ids,
                 branches=tree.branches, rates=rates, T=T, init_state=ps.init_state,
                 row_off=ps.row_off, jumps=ps.jumps, tot_time=ps.tot_time, J=J, D=D, tri_llh=tri)
        for s in SITES:
            for k, v in ref.site_debug(tree, rates, T, s).items():
                d["dbg%d_%s" % (s, k)] = np.asarray(v)
        nb = np.asarray(tree.branches) * np.linspace(0.7, 1.5, tree.n_nodes)
        ref.scale_jump_times(nb)
        sc = ref.store_paths()
        d.update(scale_branches=nb, scaled_jumps=sc.jumps)
        # a short chain from a fixed mt19937 seed (bit-exactly reproducible by the C port)
        ref.load_paths(ps)
        ref.seed(4242)
        Jm, Dm, acc, _ = ref.run_mcmc(tree, rates, T, 2, 3)
        out = ref.store_paths()
        d.update(mcmc_J=Jm, mcmc_D=Dm, mcmc_acc=acc, mcmc_init=out.init_state, mcmc_off=out.row_off,
                 mcmc_jumps=out.jumps)
        # M-step on those statistics
        llh, r2, T2, bl2 = ref.estimate_rates(Jm, Dm, rates)
        llhb, r3, br3, T3, bl3 = ref.estimate_rates_and_branches(Jm, Dm, rates, tree.branches)
        d.update(m_llh=llh, m_rates=r2, m_T=T2, m_baseline=bl2, mb_llh=llhb, mb_rates=r3,
                 mb_branches=br3, mb_T=T3)
        np.savez_compressed(os.path.join(HERE, "lib_%s.npz" % name), **d)
    # scalar known answers: transition probabilities and end-conditioned expectations
    rng = np.random.default_rng(9)
    r0 = rng.gamma(1.0, 2.0, 64) + 1e-3
    r1 = rng.gamma(1.0, 2.0, 64) + 1e-3
    t = rng.exponential(0.3, 64)
    P = np.array([ref.trans_prob_mat(a, b, c) for a, b, c in zip(r0, r1, t)])
    E = np.array([np.concatenate(ref.expectations(a, b, c)) for a, b, c in zip(r0, r1, t)])
    np.savez_compressed(os.path.join(HERE, "lib_ctmc.npz"), r0=r0, r1=r1, t=t, P=P, E=E)


if __name__ == "__main__":
    assert cpu_libs.ref_available(), "build oracle/_ref first: make -C oracle ref"
    if "--init-only" in sys.argv:        # adds the epievo_initialization fixtures, keeps the rest
        make_init()
    elif "--b-seeds-only" in sys.argv:   # adds the -b trajectories over seeds, keeps the rest
        make_cli_b_seeds()
    else:
        make_cli()
        make_cli_b_seeds()
        make_init()
        make_lib()
    print("golden vectors written to", HERE)
