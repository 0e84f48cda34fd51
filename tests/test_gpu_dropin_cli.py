"""The reference's OWN programs, compiled unchanged against the drop-in header
epievo_b200/host/SingleSiteSampler.hpp and linked with libepievo_b200.so
(integration/Makefile), run on the GPU: same flags, same files, same outputs."""
import os
import subprocess

import numpy as np
import pytest
from scipy import stats

from epievo_b200.paths import read_paths

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BUILD = os.path.join(ROOT, "integration", "_build")
CLI = os.path.join(ROOT, "tests", "golden", "cli")


def need(binary):
    p = os.path.join(BUILD, binary)
    if not os.path.exists(p):
        pytest.skip("integration/_build not built (make -C integration needs /root/reference)")
    return p


def test_est_params_histories_b200_matches_cpu_reference_over_seeds(tmp_path):
    """MCEM parameter trajectories: final parameters of `-i 3 -B 10 -L 10` over 12 seeds,
    GPU-backed binary vs the CPU reference binary (tests/golden/cli/mcem_seeds.tsv).
    Stated test: two-sample KS and Welch t per parameter, Bonferroni at alpha = 1e-3."""
    exe = need("epievo_est_params_histories_b200")
    gold = np.loadtxt(os.path.join(CLI, "mcem_seeds.tsv"), skiprows=1)[:, 1:]
    got = []
    for seed in range(101, 113):
        out, par = str(tmp_path / "o.paths"), str(tmp_path / "o.param")
        r = subprocess.run([exe, "-s", str(seed), "-i", "3", "-B", "10", "-L", "10", "-o", out, "-p", par,
                            os.path.join(CLI, "test.param"), os.path.join(CLI, "tree.nwk"),
                            os.path.join(CLI, "sim.paths")], capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
        vals = [l.split()[1:] for l in open(par)]
        got.append([float(vals[0][0]), float(vals[0][1]), float(vals[1][0]), float(vals[1][1])])
    got = np.array(got)
    for k, name in enumerate(("T00", "T11", "baseline00", "baseline11")):
        p_ks = stats.ks_2samp(gold[:, k], got[:, k]).pvalue
        p_t = stats.ttest_ind(gold[:, k], got[:, k], equal_var=False).pvalue
        assert min(p_ks, p_t) > 1e-3/8, (name, p_ks, p_t, gold[:, k].mean(), got[:, k].mean())
    # the output paths file is a valid local-paths file with the leaf data preserved
    a, b = read_paths(os.path.join(CLI, "sim.paths")), read_paths(out)
    assert (b.n_sites, b.n_nodes) == (a.n_sites, a.n_nodes) and b.node_names == a.node_names
    for v in (2, 3, 4):                                   # leaves C, D, F
        assert np.array_equal(a.end_state()[v], b.end_state()[v])


def test_reproducible_with_seed(tmp_path):
    exe = need("epievo_est_params_histories_b200")
    outs = []
    for k in range(2):
        out, par = str(tmp_path / ("o%d.paths" % k)), str(tmp_path / ("o%d.param" % k))
        r = subprocess.run([exe, "-s", "7", "-i", "2", "-B", "3", "-L", "2", "-o", out, "-p", par,
                            os.path.join(CLI, "test.param"), os.path.join(CLI, "tree.nwk"),
                            os.path.join(CLI, "sim.paths")], capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
        outs.append((open(out).read(), open(par).read()))
    assert outs[0] == outs[1]


def test_sim_pairwise_b200(tmp_path):
    """prog/epievo_sim_pairwise.cpp unchanged: its per-site Metropolis_Hastings_site loop
    becomes one GPU sweep per burn-in iteration."""
    exe = need("epievo_sim_pairwise_b200")
    rng = np.random.default_rng(4)
    n = 500
    root = (np.cumsum(rng.random(n) > 0.87) & 1)
    leaf = root ^ (rng.random(n) < 0.3)
    states = tmp_path / "pair.states"
    with open(states, "w") as fh:
        fh.write("#root\tleaf\n")
        for s in range(n):
            fh.write("%d\t%d\t%d\n" % (s, root[s], leaf[s]))
    out = str(tmp_path / "pair.paths")
    r = subprocess.run([exe, "-s", "3", "-L", "5", "-T", "1.0", "-o", out, os.path.join(CLI, "test.param"),
                        str(states)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    ps = read_paths(out)
    assert ps.n_sites == n and ps.n_nodes == 2 and ps.tot_time[1] == 1.0
    assert np.array_equal(ps.init_state[n:], root.astype(np.uint8))
    assert np.array_equal(ps.end_state()[1], leaf.astype(np.uint8))
    j = ps.jumps
    assert np.all((j > 0) & (j < 1.0))


def _em_lines(text):
    return [l.split() for l in text.splitlines() if l[:1].isdigit() and len(l.split()) == 3]


def test_epievo_initialization_b200(tmp_path):
    """prog/epievo_initialization.cpp compiled unchanged, linked with host/IndepSite_b200.cpp
    instead of IndepSite.o.  (1) The site-independent EM (:407-420) is deterministic given the
    seed (the heuristic starting paths come from the caller's mt19937 on the host): its verbose
    rate trajectory must equal the CPU reference binary's (tests/golden/cli/init.log) to the
    printed precision.  (2) The final parameters depend on the sampled histories: over 12 seeds
    they must be distributed as the reference's (init_seeds.tsv; KS and Welch t, Bonferroni
    alpha = 1e-3)."""
    exe = need("epievo_initialization_b200")
    tree, states = os.path.join(CLI, "tree.nwk"), os.path.join(CLI, "sim.states")
    out, par = str(tmp_path / "i.paths"), str(tmp_path / "i.param")
    r = subprocess.run([exe, "-v", "-s", "1", "-i", "10", "-B", "5", "-o", out, "-p", par, tree, states],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    want = _em_lines(open(os.path.join(CLI, "init.log")).read())
    got = _em_lines(r.stderr)
    assert len(got) == len(want) == 11
    for a, b in zip(got, want):
        assert a[0] == b[0]
        assert np.allclose([float(a[1]), float(a[2])], [float(b[1]), float(b[2])], rtol=2e-6, atol=0)
    ps = read_paths(out)                                  # a valid local-paths file over the leaf data
    leaves = np.loadtxt(states, skiprows=1, dtype=int)[:, 1:]
    names = open(states).readline().lstrip("#").split()
    for v, nm in enumerate(ps.node_names):
        if nm in ("C", "D", "F"):
            assert np.array_equal(ps.end_state()[v], leaves[:, names.index(nm)])
    gold = np.loadtxt(os.path.join(CLI, "init_seeds.tsv"), skiprows=1)[:, 1:]
    rows = []
    for seed in range(201, 213):
        r = subprocess.run([exe, "-s", str(seed), "-i", "10", "-B", "5", "-o", out, "-p", par, tree, states],
                           capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
        vals = [l.split()[1:] for l in open(par)]
        rows.append([float(vals[0][0]), float(vals[0][1]), float(vals[1][0]), float(vals[1][1])])
    rows = np.array(rows)
    for k, name in enumerate(("T00", "T11", "baseline00", "baseline11")):
        p_ks = stats.ks_2samp(gold[:, k], rows[:, k]).pvalue
        p_t = stats.ttest_ind(gold[:, k], rows[:, k], equal_var=False).pvalue
        assert min(p_ks, p_t) > 1e-3/8, (name, p_ks, p_t, gold[:, k].mean(), rows[:, k].mean())


def test_est_complete_b200_writes_the_reference_parameter_file(tmp_path):
    """Config 3's program: `epievo_est_complete` (J/D pass over every path + M-step) compiled unchanged and
    linked with host/ParamEstimation_b200.cpp, which replaces get_sufficient_statistics / estimate_rates on
    paths at link time.  No RNG on this path: the parameter file must be the one the CPU reference binary wrote
    (tests/golden/cli/est_complete.param) -- identical text, or (D goes through 2^-40 fixed point on the
    device) every number within 1e-9 relative."""
    exe = need("epievo_est_complete_b200")
    par = str(tmp_path / "c.param")
    r = subprocess.run([exe, "-v", "-o", par, os.path.join(CLI, "test.param"), os.path.join(CLI, "tree.nwk"),
                        os.path.join(CLI, "sim.paths")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    assert "[COMPUTING INITIAL SUFFICIENT STATISTICS]" in r.stderr          # the replaced function ran
    got, want = open(par).read(), open(os.path.join(CLI, "est_complete.param")).read()
    if got != want:
        g = [float(x) for l in got.splitlines() for x in l.split()[1:]]
        w = [float(x) for l in want.splitlines() for x in l.split()[1:]]
        assert [l.split()[0] for l in got.splitlines()] == [l.split()[0] for l in want.splitlines()]
        assert np.allclose(g, w, rtol=1e-9, atol=0), (got, want)
