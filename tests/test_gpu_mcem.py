"""Device-resident MCEM loop (epv_mcem_iteration = the body of the loop at
prog/epievo_est_params_histories.cpp:239-264) against the CPU reference binary's parameter
trajectories over seeds (tests/golden/cli/mcem_seeds.tsv: -i 3 -B 10 -L 10 on the
walk-through data).  Stated test: KS and Welch t per parameter, Bonferroni alpha = 1e-3."""
import os

import numpy as np
import pytest
from scipy import stats

import cpu_libs
from epievo_b200.paths import read_paths, read_tree
from test_golden_oracle import CLI, read_param_file

pytestmark = pytest.mark.gpu


def run_mcem(ctx, seed, optimize_branches=False, iterations=3):
    tree = read_tree(os.path.join(CLI, "tree.nwk"))
    ps = read_paths(os.path.join(CLI, "sim.paths"))
    rates, T, _, _ = cpu_libs.oracle().model_from_params(*read_param_file(os.path.join(CLI, "test.param")))
    ctx.set_tree(tree); ctx.set_model(rates, T); ctx.upload_paths(ps)
    branches = np.array(tree.branches)
    sweep = 0
    for _ in range(iterations):
        rates, T, branches, bl, acc, llh = ctx.mcem_iteration(10, 10, seed, sweep, rates, T, branches,
                                                              optimize_branches)
        sweep += 20
        assert 0.9 < acc <= 1.0 and np.isfinite(llh)
    return T, bl, branches, ctx.download_paths(), ps


def test_mcem_parameter_trajectories_match_cpu_reference(ctx):
    gold = np.loadtxt(os.path.join(CLI, "mcem_seeds.tsv"), skiprows=1)[:, 1:]
    got = []
    for seed in range(301, 313):
        T, bl, _, out, ps = run_mcem(ctx, seed)
        got.append([T[0], T[3], bl[0], bl[3]])
    got = np.array(got)
    for k, name in enumerate(("T00", "T11", "baseline00", "baseline11")):
        p_ks = stats.ks_2samp(gold[:, k], got[:, k]).pvalue
        p_t = stats.ttest_ind(gold[:, k], got[:, k], equal_var=False).pvalue
        assert min(p_ks, p_t) > 1e-3/8, (name, p_ks, p_t, gold[:, k].mean(), got[:, k].mean())
    # the paths came back rescaled to the final branch lengths, leaf data untouched
    for v in (2, 3, 4):
        assert np.array_equal(out.end_state()[v], ps.end_state()[v])


def test_mcem_with_branch_updates_trajectories_match_cpu_reference(ctx):
    """The -b path (estimate_rates_and_branches + scale_jump_times every iteration,
    prog/epievo_est_params_histories.cpp:259-264): final parameters AND branch lengths after `-b -i 3
    -B 10 -L 10` over 12 seeds, device loop vs the CPU reference binary (tests/golden/cli/mcem_b_seeds.tsv).
    Stated test: KS and Welch t per quantity (4 parameters + 4 branch lengths), Bonferroni alpha = 1e-3."""
    gold = np.loadtxt(os.path.join(CLI, "mcem_b_seeds.tsv"), skiprows=1)[:, 1:]
    got = []
    for seed in range(401, 413):
        T, bl, branches, out, ps = run_mcem(ctx, seed, optimize_branches=True)
        assert np.allclose(out.tot_time[1:], branches[1:])
        got.append([T[0], T[3], bl[0], bl[3]] + list(branches[1:]))
    got = np.array(got)
    names = ("T00", "T11", "baseline00", "baseline11", "branch1", "branch2", "branch3", "branch4")
    for k, name in enumerate(names):
        p_ks = stats.ks_2samp(gold[:, k], got[:, k]).pvalue
        p_t = stats.ttest_ind(gold[:, k], got[:, k], equal_var=False).pvalue
        assert min(p_ks, p_t) > 1e-3/16, (name, p_ks, p_t, gold[:, k].mean(), got[:, k].mean())


def test_mcem_with_branch_updates(ctx):
    """One -b run, structure: paths rescaled to the final branch lengths, every jump inside its branch, and
    the branch lengths inside the spread (mean +- 5 sd over 12 seeds) of the CPU reference binary's."""
    T, bl, branches, out, ps = run_mcem(ctx, 5, optimize_branches=True, iterations=3)
    gold = np.loadtxt(os.path.join(CLI, "mcem_b_seeds.tsv"), skiprows=1)[:, 5:9]
    assert np.allclose(out.tot_time[1:], branches[1:])
    assert np.all(np.abs(branches[1:] - gold.mean(0)) < 5*gold.std(0, ddof=1)), (branches[1:], gold.mean(0), gold.std(0, ddof=1))
    j = out.jumps
    t = np.repeat(out.tot_time, out.n_sites)[np.repeat(np.arange(out.n_nodes*out.n_sites), out.row_len().ravel())]
    assert np.all((j > 0) & (j < t))
