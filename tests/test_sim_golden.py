"""The committed statistics of the reference's forward simulator (tests/golden/sim_stats.json,
made by tests/golden/make_sim_golden.py from the reference binaries) against what the model
fixes exactly: every interior site spends the whole branch in some triplet, J counts the jumps,
the rates are scaled to one change per site and unit time, and the root chain is stationary."""
import json
import os

import numpy as np

from epievo_b200.paths import Tree
from synth import ref_test_tree

GOLDEN = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "sim_stats.json")))
TREES = {"testtree": ref_test_tree, "pair_T1": lambda: Tree.pair(1.0)}


def test_golden_simulator_statistics_are_self_consistent(model):
    rates, T = model
    pi1 = (1.0 - T[0]) / (2.0 - T[3] - T[0])
    for case, g in GOLDEN.items():
        tree = TREES[case]()
        br = np.asarray(tree.branches, np.float64)
        J, D = np.asarray(g["J"]["mean"]), np.asarray(g["D"]["mean"])
        assert J.shape == (tree.n_nodes, 8) and np.all(J[0] == 0) and np.all(D[0] == 0)
        # time is conserved per site: sum_t D[b][t] == branch length (exact up to rounding)
        assert np.allclose(D[1:].sum(1), br[1:], rtol=1e-9)
        # J is the number of jumps
        assert np.allclose(J.sum(1), g["jumps_per_site"]["mean"], rtol=1e-9)
        # one change per site and unit time at stationarity (EpiEvoModel.cpp:372-377): within 2 %
        assert np.allclose(np.asarray(g["jumps_per_site"]["mean"])[1:], br[1:], rtol=2e-2)
        # the fraction of 1s stays at the horizontal chain's stationary value down the tree
        f0, f1 = np.asarray(g["frac1_start"]["mean"])[1:], np.asarray(g["frac1_end"]["mean"])[1:]
        assert np.all(np.abs(f0 - pi1) < 5e-3) and np.all(np.abs(f1 - pi1) < 5e-3)
        # J / D estimates the triplet rates (the M-step's fixed point), branch by branch: within 5 %
        for b in range(1, tree.n_nodes):
            assert np.allclose(J[b] / D[b], rates, rtol=5e-2), (case, b, J[b] / D[b], rates)


def test_thinning_algorithm_matches_reference_simulator(model):
    """The numpy restatement of the device algorithm (oracle/sim_oracle.py: thinning + decisions in
    local-minimum order) against the reference simulator's statistics: same z-test as the device
    test (|z| < 4.5 on every per-site statistic), 8 blocks of 50 000 sites per case."""
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(__file__)), "oracle"))
    import sim_oracle
    from sim_stats import stats_of
    rates, T = model
    for case, gold in GOLDEN.items():
        tree = TREES[case]()
        per = [stats_of(sim_oracle.simulate_paths(tree, 50_000, rates, T, seed=100 + b)) for b in range(8)]
        for key, g in gold.items():
            if not isinstance(g, dict):
                continue
            a = np.stack([np.asarray(p[key], np.float64) for p in per])
            mean, sem = a.mean(0), a.std(0, ddof=1) / np.sqrt(len(per))
            gm, gs = np.asarray(g["mean"]), np.asarray(g["sem"])
            den = np.sqrt(sem**2 + gs**2)
            z = np.where(den > 0, (mean - gm) / np.where(den > 0, den, 1.0), 0.0)
            assert np.abs(z).max() < 4.5, (case, key, z)
