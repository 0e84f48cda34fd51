"""Exact-posterior validation of a single-site sampler (the design of the reference's
harness, /root/reference/src/harnesses/MCMC_test.cpp): on a tiny single-branch problem the
posterior over histories given the root and leaf sequences can be sampled EXACTLY by forward
simulation of the whole sequence + rejection on the leaf sequence
(end_cond_forward_simulation, MCMC_test.cpp:192-216, loop :366-379).  The per-draw
sufficient statistics J[8], D[8] of exact draws and of well-separated sampler draws must then
have the same distribution.

The exact draws (and the reference CPU sampler's draws, used for attribution) come from the
UNMODIFIED reference harness, oracle/_ref/bin/MCMC_test, built by oracle/Makefile; nothing
here reads /root/reference at run time.

Stated test, per case, family-wise alpha = 1e-3 (Bonferroni over the 16 statistics):
  * J_t (integer jump count in context t over the interior sites): two-sample chi-square on
    the histogram of J_t, cells merged from the top until every pooled cell holds >= 10 draws;
  * D_t (time spent in context t): two-sample Kolmogorov-Smirnov.
Statistics that are constant in both arms must be equal.
"""
from __future__ import annotations

import os
import re
import subprocess
from typing import Dict, List, Tuple

import numpy as np
from scipy import stats

from epievo_b200.paths import PathSet, Tree

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_MCMC_TEST = os.path.join(ROOT, "oracle", "_ref", "bin", "MCMC_test")
B200_MCMC_TEST = os.path.join(ROOT, "integration", "_build", "MCMC_test_b200")
PARAM = os.path.join(ROOT, "tests", "golden", "cli", "test.param")

ALPHA = 1e-3


def read_stats(path: str) -> np.ndarray:
    """[N, 16]: J_000..J_111, D_000..D_111 per sample (write_statistics, MCMC_test.cpp:253-262)."""
    return np.loadtxt(path, skiprows=1)[:, 1:17]


def run_harness(exe: str, n_sites: int, evo_time: float, n_draws: int, thin: int, seed: int,
                prefix: str) -> Tuple[np.ndarray, np.ndarray, np.ndarray, np.ndarray]:
    """Runs MCMC_test (reference or drop-in build).  Returns root sequence, leaf sequence,
    the exact forward-rejection draws and the sampler's draws (thin burn-in sweeps + one
    recorded sweep per draw, MCMC_test.cpp:394-403)."""
    r = subprocess.run([exe, "-n", str(n_sites), "-T", repr(evo_time), "-B", str(n_draws), "-L", str(thin),
                        "-s", str(seed), "-v", "-o", prefix, PARAM], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-2000:]
    root = re.search(r"ROOT SEQ: ([01]+)", r.stderr).group(1)
    leaf = re.search(r"END SEQ: ([01]+)", r.stderr).group(1)
    as_arr = lambda s: np.array([int(c) for c in s], np.uint8)
    return as_arr(root), as_arr(leaf), read_stats(prefix + ".forward"), read_stats(prefix + ".mcmc")


def initial_paths(root: np.ndarray, leaf: np.ndarray, evo_time: float, seed: int) -> Tuple[Tree, PathSet]:
    """initialize_paths (MCMC_test.cpp:219-237): one uniformly placed jump where root != leaf."""
    rng = np.random.default_rng(seed)
    n = len(root)
    rows = [[[] for _ in range(n)], [[rng.random()*evo_time] if root[s] != leaf[s] else [] for s in range(n)]]
    tree = Tree.pair(evo_time)
    ps = PathSet.from_rows(n, 2, [np.zeros(n, np.uint8), root], rows, np.array([0.0, evo_time]))
    return tree, ps


def round6(x: np.ndarray) -> np.ndarray:
    """Values as the harness prints them (ostream default precision, 6 significant digits), so that
    an in-process arm and a harness arm put their atoms (D_t = 0, = T, ...) at identical values."""
    return np.array([float("%.6g" % v) for v in np.asarray(x, np.float64).ravel()]).reshape(np.shape(x))


def chi2_two_sample(a: np.ndarray, b: np.ndarray) -> float:
    """p-value of the two-sample chi-square test on integer-valued samples (cells merged from
    the top until every pooled cell holds >= 10 draws)."""
    a = np.rint(a).astype(np.int64); b = np.rint(b).astype(np.int64)
    top = int(max(a.max(), b.max()))
    ca = np.bincount(a, minlength=top + 1).astype(np.float64)
    cb = np.bincount(b, minlength=top + 1).astype(np.float64)
    # merge sparse cells: walk from the top, fold into the cell below
    cells_a: List[float] = []; cells_b: List[float] = []
    acc_a = acc_b = 0.0
    for k in range(top, -1, -1):
        acc_a += ca[k]; acc_b += cb[k]
        if acc_a + acc_b >= 10:
            cells_a.append(acc_a); cells_b.append(acc_b); acc_a = acc_b = 0.0
    if acc_a + acc_b > 0:
        if cells_a:
            cells_a[-1] += acc_a; cells_b[-1] += acc_b
        else:
            cells_a.append(acc_a); cells_b.append(acc_b)
    # drop empty pooled cells (values nobody took)
    tab = np.array([[x, y] for x, y in zip(cells_a, cells_b) if x + y > 0]).T
    if tab.shape[1] < 2:
        return 1.0
    return float(stats.chi2_contingency(tab, correction=False)[1])


def compare(exact: np.ndarray, chain: np.ndarray) -> Dict[str, float]:
    """p-values of the stated tests, keyed J_000.. / D_000..; constant statistics must agree."""
    out: Dict[str, float] = {}
    for t in range(8):
        name = format(t, "03b")
        ja, jb = exact[:, t], chain[:, t]
        if ja.min() == ja.max() and jb.min() == jb.max():
            assert ja[0] == jb[0], ("J_" + name, ja[0], jb[0])
        else:
            out["J_" + name] = chi2_two_sample(ja, jb)
        da, db = round6(exact[:, 8 + t]), round6(chain[:, 8 + t])
        if da.min() == da.max() and db.min() == db.max():
            assert abs(da[0] - db[0]) <= 1e-6*max(1.0, abs(da[0])), ("D_" + name, da[0], db[0])   # text output: 6 digits
        else:
            out["D_" + name] = float(stats.ks_2samp(da, db).pvalue)
    return out


def assert_same_law(exact: np.ndarray, chain: np.ndarray, label: str) -> Dict[str, float]:
    p = compare(exact, chain)
    thr = ALPHA/16
    bad = {k: v for k, v in p.items() if v < thr}
    assert not bad, "%s: rejected at family-wise alpha=%g: %s (means exact %s vs chain %s)" % (
        label, ALPHA, bad, np.round(exact.mean(0), 4), np.round(chain.mean(0), 4))
    return p
