"""Pins the C restatement (oracle/liboracle.so) against the UNMODIFIED reference build
(oracle/_ref/libepievo_ref.so, compiled from /root/reference by oracle/Makefile):
every entry point on identical inputs, bit-exact -- including whole run_mcmc chains
driven by the same mt19937 seed.  Skipped where the reference build is absent."""
import numpy as np
import pytest

import cpu_libs
from epievo_b200.paths import Tree
from epievo_b200.synthetic import simulate_paths
from synth import TEST_PARAMS, ref_test_tree, synth_paths

pytestmark = pytest.mark.skipif(not cpu_libs.ref_available(),
                                reason="oracle/_ref not built (needs /root/reference)")


@pytest.fixture(scope="module")
def libs():
    return cpu_libs.oracle(), cpu_libs.ref()


TREES = {"testtree": ref_test_tree, "pair": lambda: Tree.pair(1.0),
         "bal8": lambda: Tree.balanced(8, 0.05), "bal32": lambda: Tree.balanced(32, 0.05)}


def test_model_arithmetic(libs):
    o, r = libs
    for params in (TEST_PARAMS, (0.7, 0.6, -0.3, -1.1), (0.95, 0.9, -2.0, -0.5)):
        for scale in (0, 1):
            a = o.model_from_params(*params, scale)
            b = r.model_from_params(*params, scale)
            for x, y in zip(a, b):
                assert np.array_equal(x, y)
        rates = o.model_from_params(*params, 1)[0]
        assert o.rate_scaling_factor(rates) == r.rate_scaling_factor(rates)
        for x, y in zip(o.model_rebuild(rates), r.model_rebuild(rates)):
            assert np.array_equal(x, y)


def test_transition_probs_and_expectations(libs):
    o, r = libs
    rng = np.random.default_rng(3)
    for _ in range(200):
        r0, r1 = rng.gamma(1.0, 2.0, 2) + 1e-3
        t = rng.exponential(0.3)
        assert np.array_equal(o.trans_prob_mat(r0, r1, t), r.trans_prob_mat(r0, r1, t))
        for a in (0, 1):
            for b in (0, 1):
                assert o.trans_prob(r0, r1, t, a, b) == r.trans_prob(r0, r1, t, a, b)
        for x, y in zip(o.expectations(r0, r1, t), r.expectations(r0, r1, t)):
            assert np.array_equal(x, y)


def test_segment_merge_and_triple_stats(libs):
    o, r = libs
    rates = o.model_from_params(*TEST_PARAMS)[0]
    rng = np.random.default_rng(5)
    for it in range(300):
        T = 0.5
        rows = []
        for _ in range(3):
            k = rng.integers(0, 5)
            rows.append((int(rng.integers(0, 2)), np.sort(rng.random(k))*T))
        if it % 7 == 0 and len(rows[0][1]) and len(rows[2][1]):   # exact ties
            rows[2][1][0] = rows[0][1][0]
            rows[2] = (rows[2][0], np.sort(rows[2][1]))
        if it % 11 == 0 and len(rows[1][1]) and len(rows[2][1]):
            rows[1][1][0] = rows[2][1][0]
            rows[1] = (rows[1][0], np.sort(rows[1][1]))
        a = o.segment_info(rates, rows[0][0], rows[0][1], rows[2][0], rows[2][1], T)
        b = r.segment_info(rates, rows[0][0], rows[0][1], rows[2][0], rows[2][1], T)
        for x, y in zip(a, b):
            assert np.array_equal(x, y)
        Ja, Da = o.add_suff_stats(rows[0], rows[1], rows[2], T)
        Jb, Db = r.add_suff_stats(rows[0], rows[1], rows[2], T)
        assert np.array_equal(Ja, Jb) and np.array_equal(Da, Db)


@pytest.mark.parametrize("case", sorted(TREES))
def test_whole_data_passes(libs, case):
    o, r = libs
    tree = TREES[case]()
    rates, T, _, _ = o.model_from_params(*TEST_PARAMS)
    ps = synth_paths(tree, 400, 17, density=4.0)
    o.load_paths(ps); r.load_paths(ps)
    assert o.store_paths().same_as(ps) and r.store_paths().same_as(ps)
    for x, y in zip(o.get_suff_stats(), r.get_suff_stats()):
        assert np.array_equal(x, y)
    assert np.array_equal(o.tri_llh(rates, T), r.tri_llh(rates, T))
    for site in (1, 2, 57, 200, 398):
        a, b = o.site_debug(tree, rates, T, site), r.site_debug(tree, rates, T, site)
        for k in a:
            assert np.array_equal(a[k], b[k]), (site, k)
    nb = np.asarray(tree.branches)*np.linspace(0.6, 1.7, tree.n_nodes)
    o.scale_jump_times(nb); r.scale_jump_times(nb)
    assert o.store_paths().same_as(r.store_paths())


@pytest.mark.parametrize("case", ["testtree", "pair", "bal8"])
def test_run_mcmc_chain_bit_exact(libs, case):
    """Same mt19937 seed => identical chains: J, D, acceptance rate and final paths."""
    o, r = libs
    tree = TREES[case]()
    rates, T, _, _ = o.model_from_params(*TEST_PARAMS)
    ps = simulate_paths(tree, 300, rates, T, 23)
    o.load_paths(ps); r.load_paths(ps)
    o.seed(99); r.seed(99)
    Jo, Do, ao, _ = o.run_mcmc(tree, rates, T, 3, 4)
    Jr, Dr, ar, _ = r.run_mcmc(tree, rates, T, 3, 4)
    assert np.array_equal(Jo, Jr) and np.array_equal(Do, Dr) and ao == ar
    assert o.store_paths().same_as(r.store_paths())


def test_end_cond_sampler_stream(libs):
    o, r = libs
    o.seed(5); r.seed(5)
    for a in (0, 1):
        for b in (0, 1):
            for _ in range(50):
                assert np.array_equal(o.end_cond_sample(1.5, 0.5, a, b, 1.0),
                                      r.end_cond_sample(1.5, 0.5, a, b, 1.0))


def test_m_step(libs):
    o, r = libs
    tree = ref_test_tree()
    rates, T, _, _ = o.model_from_params(*TEST_PARAMS)
    ps = simulate_paths(tree, 5000, rates, T, 31)
    r.load_paths(ps)
    J, D = r.get_suff_stats()
    start = o.model_from_params(0.8, 0.8, -0.5, -1.0)[0]
    a = o.estimate_rates(J, D, start); b = r.estimate_rates(J, D, start)
    assert a[0] == b[0]
    for x, y in zip(a[1:], b[1:]):
        assert np.array_equal(x, y)
    a = o.estimate_rates_and_branches(J, D, start, tree.branches)
    b = r.estimate_rates_and_branches(J, D, start, tree.branches)
    assert a[0] == b[0]
    for x, y in zip(a[1:], b[1:]):
        assert np.array_equal(x, y)
    a = o.set_one_change(start, tree.branches); b = r.set_one_change(start, tree.branches)
    for x, y in zip(a, b):
        assert np.array_equal(x, y)
