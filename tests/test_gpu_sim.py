"""epv_simulate (forward simulation on the device) against the reference's
epievo_sim.  The two simulators share a law, not a random stream: structure is checked exactly,
the law through the per-site statistics of tests/golden/sim_stats.json (reference binaries,
tests/golden/make_sim_golden.py) with a z-test whose threshold is stated below."""
import json
import os

import numpy as np
import pytest

from epievo_b200.paths import Tree
from sim_stats import stats_of
from synth import ref_test_tree

pytestmark = pytest.mark.gpu

GOLDEN = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "sim_stats.json")))
CASES = {"testtree": ref_test_tree, "pair_T1": lambda: Tree.pair(1.0)}


def simulate(ctx, model, tree, n, seed, root=None):
    ctx.set_tree(tree); ctx.set_model(*model)
    return ctx.simulate(n, seed, tree.branches, root_states=root, node_names=tree.node_names)


@pytest.mark.parametrize("case,n", [("testtree", 5000), ("pair_T1", 3000), ("testtree", 3), ("testtree", 513)])
def test_structure(ctx, model, case, n):
    tree = CASES[case]()
    ps = simulate(ctx, model, tree, n, seed=3)
    assert ps.n_sites == n and ps.n_nodes == tree.n_nodes
    lens = ps.row_len()
    init = ps.init_state.reshape(ps.n_nodes, n).astype(np.int64)
    end = init ^ (lens & 1)
    assert lens[0].sum() == 0                                     # root rows are empty
    assert lens[:, 0].sum() == 0 and lens[:, -1].sum() == 0       # the end sites never change
    for v in range(1, tree.n_nodes):
        p = int(tree.parent_ids[v])
        if p > 0:
            assert np.array_equal(init[v], end[p])                # a branch starts where its parent ended
        else:
            assert np.array_equal(init[v], init[1])               # children of the root share the root sequence
        T = float(tree.branches[v])
        for s in range(0, n, max(1, n // 400)):
            r = ps.row(v, s)
            assert np.all(r > 0.0) and np.all(r < T) and np.all(np.diff(r) > 0.0)
    # the stream depends on the seed only
    again = simulate(ctx, model, tree, n, seed=3)
    assert again.same_as(ps)
    if n > 100:
        assert not simulate(ctx, model, tree, n, seed=4).same_as(ps)


def test_given_root_and_sampler_accepts_it(ctx, model):
    tree = ref_test_tree()
    n = 4000
    root = (np.arange(n) // 7 % 2).astype(np.uint8)
    ps = simulate(ctx, model, tree, n, seed=11, root=root)
    assert np.array_equal(ps.init_state.reshape(tree.n_nodes, n)[1], root)
    ctx.upload_paths(ps)
    assert ctx.download_paths().same_as(ps)
    ctx.reset()
    J, D, acc = ctx.run_mcmc(1, 1, seed=5)
    assert 0.0 < acc <= 1.0 and np.all(np.isfinite(D))


@pytest.mark.parametrize("case", sorted(CASES))
def test_law_matches_reference_simulator(ctx, model, case):
    """z = (ours - reference) / sqrt(sem_ours^2 + sem_ref^2) for every per-site statistic (J and D per
    branch and triplet, jumps per branch, fraction of 1s at the ends of every branch); ours from 12
    blocks of 50 000 sites.  ~170 statistics: |z| < 4.5 everywhere (two-sided p = 7e-6 each)."""
    tree = CASES[case]()
    gold = GOLDEN[case]
    block, n_blocks = 50_000, 12
    ps = simulate(ctx, model, tree, block * n_blocks, seed=20240)
    per = [stats_of(ps.site_slice(b * block, (b + 1) * block)) for b in range(n_blocks)]
    worst = 0.0
    for key, g in gold.items():
        if not isinstance(g, dict):
            continue
        a = np.stack([np.asarray(p[key], np.float64) for p in per])
        mean, sem = a.mean(0), a.std(0, ddof=1) / np.sqrt(n_blocks)
        gm, gs = np.asarray(g["mean"]), np.asarray(g["sem"])
        den = np.sqrt(sem**2 + gs**2)
        z = np.where(den > 0, (mean - gm) / np.where(den > 0, den, 1.0), 0.0)
        assert np.all(np.abs(mean - gm)[den == 0] < 1e-12), key     # exact zeros (root rows)
        worst = max(worst, float(np.abs(z).max()))
        assert np.abs(z).max() < 4.5, (key, z)
    assert worst > 0.0
