"""The exact-posterior test (tests/exact_posterior.py) applied to the REFERENCE's CPU sampler:
pins the test itself (the reference's SingleSiteSampler must pass it) and gives the attribution
arm of tests/test_gpu_exact_posterior.py."""
import os

import pytest

import exact_posterior as ep

pytestmark = pytest.mark.skipif(not os.path.exists(ep.REF_MCMC_TEST), reason="oracle/_ref not built")


@pytest.mark.parametrize("n,T,seed", [(5, 0.1, 7), (5, 1.0, 7), (8, 1.0, 7), (12, 0.1, 3), (12, 1.0, 5)])
def test_reference_sampler_matches_exact_posterior(tmp_path, n, T, seed):
    root, leaf, exact, chain = ep.run_harness(ep.REF_MCMC_TEST, n, T, 20000, 10, seed, str(tmp_path / "ref"))
    assert root[0] == leaf[0] and root[-1] == leaf[-1]       # the end sites never change (TripletSampler.cpp:58-60)
    ep.assert_same_law(exact, chain, "reference sampler n=%d T=%g" % (n, T))


def test_the_test_has_power(tmp_path):
    """A sampler with a slightly wrong law must be rejected: the exact draws of evolutionary time
    1.0 against chain draws of time 1.05 on the same root/leaf pair."""
    _, _, exact, _ = ep.run_harness(ep.REF_MCMC_TEST, 8, 1.0, 20000, 1, 7, str(tmp_path / "a"))
    _, _, other, _ = ep.run_harness(ep.REF_MCMC_TEST, 8, 1.05, 20000, 1, 7, str(tmp_path / "b"))
    p = ep.compare(exact, other)
    assert min(p.values()) < ep.ALPHA/16


# ---- trees: exact draws from tests/exact_tree.py (numpy forward simulation + rejection) --------------
import numpy as np

import exact_tree as et
from epievo_b200.paths import Tree, parse_newick

TREES = {
    # the reference's test tree with every branch x5 (so that histories hold jumps), and a cherry
    "testtree_x5": "((C:0.15,D:0.3)E:0.1,F:0.5)G:0.0;",
    "cherry": "(A:0.4,B:0.7)R:0.0;",
}


def tree_case(name, n_sites, model, seed):
    tree = parse_newick(TREES[name])
    rates, T = model
    root, leaves = et.make_problem(tree, n_sites, rates, T, seed)
    return tree, root, leaves


def compare_tree(exact, chain, label):
    """exact, chain: [N, n_nodes, 16].  The stated test of exact_posterior.compare on every branch,
    Bonferroni over all branches (family-wise alpha = 1e-3)."""
    n_br = exact.shape[1] - 1
    worst = 1.0
    for v in range(1, exact.shape[1]):
        p = ep.compare(exact[:, v], chain[:, v])
        bad = {k: x for k, x in p.items() if x < ep.ALPHA/(16*n_br)}
        assert not bad, "%s branch %d: %s (means exact %s vs chain %s)" % (
            label, v, bad, np.round(exact[:, v].mean(0), 4), np.round(chain[:, v].mean(0), 4))
        worst = min(worst, min(p.values()) if p else 1.0)
    return worst


def test_numpy_exact_sampler_matches_reference_harness(tmp_path, model):
    """Pins tests/exact_tree.py: on one branch its draws must have the law of the reference
    harness's own forward-rejection draws (same root / leaf pair)."""
    root, leaf, exact_ref, _ = ep.run_harness(ep.REF_MCMC_TEST, 8, 1.0, 20000, 1, 7, str(tmp_path / "r"))
    mine, acc = et.exact_draws(Tree.pair(1.0), root, {1: leaf}, model[0], 20000, seed=5)
    ep.assert_same_law(exact_ref, mine[:, 1], "numpy exact sampler vs reference harness (acceptance %.3g)" % acc)


def cpu_chain(cpu, tree, model, ps, n_draws, thin, seed):
    rates, T = model
    cpu.load_paths(ps); cpu.seed(seed)
    out = np.zeros((n_draws, tree.n_nodes, 16))
    for k in range(n_draws):
        J, D, _, _ = cpu.run_mcmc(tree, rates, T, thin, 1, reset=(k == 0))
        out[k, :, :8] = J; out[k, :, 8:] = D
    return out


@pytest.mark.parametrize("name,n,seed", [("testtree_x5", 6, 6), ("cherry", 7, 4)])
def test_reference_sampler_matches_exact_posterior_on_a_tree(cpu, model, name, n, seed):
    tree, root, leaves = tree_case(name, n, model, seed)
    exact, acc = et.exact_draws(tree, root, leaves, model[0], 20000, seed=100 + seed)
    chain = cpu_chain(cpu, tree, model, et.initial_paths(tree, root, leaves, seed), 20000, 10, 77)
    worst = compare_tree(exact, chain, "reference sampler on " + name)
    print("%s: acceptance of the exact sampler %.3g, smallest p %.3g" % (name, acc, worst))
