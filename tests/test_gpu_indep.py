"""Site-independent E-step of epievo_initialization on the GPU against the
checker (oracle/indep_oracle.py, pinned to IndepSite.cpp by tests/test_indep_cpu.py; the
unmodified reference itself when oracle/_ref is present):
* epv_indep_expectation / epv_indep_suffstats: J exact, expectations within 1e-9 relative;
* the deterministic EM loop of epievo_initialization.cpp:407-420 reproduces the reference's
  rate trajectory;
* epv_indep_update_paths: the new histories keep root and leaf states, are valid paths, their
  J/D average to the analytic expectation (z-test, Bonferroni alpha = 1e-3) and match the
  reference sampler's (Welch t, same alpha); shards reproduce the single-GPU result bit for bit."""
import os
import sys

import numpy as np
import pytest
from scipy import stats

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "oracle"))
import cpu_libs  # noqa: E402
import indep_oracle as IO  # noqa: E402
from epievo_b200 import _lib as L  # noqa: E402
from epievo_b200 import sharded as S  # noqa: E402
from epievo_b200.device import Context  # noqa: E402
from epievo_b200.paths import Tree  # noqa: E402
from epievo_b200.synthetic import simulate_paths  # noqa: E402
from synth import ref_test_tree  # noqa: E402

pytestmark = pytest.mark.gpu

TREES = {"testtree": ref_test_tree, "pair": lambda: Tree.pair(1.0),
         "bal8": lambda: Tree.balanced(8, 0.05), "bal32": lambda: Tree.balanced(32, 0.3),
         "bal64": lambda: Tree.balanced(64, 0.1)}


def _ctx(tree, ps, model):
    c = Context(0)
    c.set_tree(tree); c.set_model(*model); c.upload_paths(ps)
    return c


@pytest.mark.parametrize("case,n", [("testtree", 3), ("testtree", 1000), ("pair", 777), ("bal8", 2049),
                                    ("bal32", 300), ("bal64", 65)])
@pytest.mark.parametrize("rates", [(1.3, 0.7), (0.2, 2.5)])
def test_expectation_and_suffstats(model, case, n, rates):
    tree = TREES[case]()
    ps = simulate_paths(tree, n, model[0], model[1], seed=21)
    c = _ctx(tree, ps, model)
    try:
        J, D = c.indep_expectation(rates)
        Jo, Do = IO.expectation_sufficient_statistics(tree, ps, rates)
        assert np.allclose(J, Jo, rtol=1e-9, atol=1e-12) and np.allclose(D, Do, rtol=1e-9, atol=1e-12)
        J, D = c.indep_suffstats()
        Jo, Do = IO.compute_sufficient_statistics(ps)
        assert np.array_equal(J, Jo) and np.allclose(D, Do, rtol=1e-9, atol=0)
        # total time is conserved
        assert np.allclose(D[1:].sum(1), n * np.asarray(tree.branches)[1:], rtol=1e-9)
    finally:
        c.close()


def test_em_loop_matches_reference(model):
    """epievo_initialization.cpp:401-420: J, D from the initial paths, then
    (estimate_rates_indep, expectation_sufficient_statistics) x iterations."""
    tree = ref_test_tree()
    ps = simulate_paths(tree, 5000, model[0], model[1], seed=8)
    lib = L.load()
    c = _ctx(tree, ps, model)
    try:
        J, D = c.indep_suffstats()
        J, D = J / ps.n_sites, D / ps.n_sites
        rates = np.zeros(2)
        traj = []
        for _ in range(6):
            lib.epv_indep_estimate_rates(tree.n_nodes, J.ctypes.data_as(L.dp), D.ctypes.data_as(L.dp),
                                         rates.ctypes.data_as(L.dp))
            J, D = c.indep_expectation(rates)
            J, D = np.ascontiguousarray(J), np.ascontiguousarray(D)
            traj.append(rates.copy())
    finally:
        c.close()
    # the same loop through the checker (the reference build when present)
    if cpu_libs.ref_available():
        ref = cpu_libs.ref()
        ref.indep_load(ps)
        Jr, Dr = ref.indep_suffstats()
        r = np.zeros(2)
        for k in range(6):
            r, _ = ref.indep_estimate(tree, Jr, Dr, r, False)
            Jr, Dr = ref.indep_expectation(tree, r)
            assert np.allclose(traj[k], r, rtol=1e-9), (k, traj[k], r)
    else:
        Jr, Dr = IO.compute_sufficient_statistics(ps)
        Jr, Dr = Jr / ps.n_sites, Dr / ps.n_sites
        r = np.zeros(2)
        for k in range(6):
            r = np.array([max(Jr[1:, 0].sum() / Dr[1:, 0].sum(), 10e-6), max(Jr[1:, 1].sum() / Dr[1:, 1].sum(), 10e-6)])
            Jr, Dr = IO.expectation_sufficient_statistics(tree, ps, r)
            assert np.allclose(traj[k], r, rtol=1e-9)


@pytest.mark.parametrize("case,n", [("testtree", 4000), ("pair", 3000), ("bal8", 3000)])
def test_update_paths_valid_and_unbiased(model, case, n):
    tree = TREES[case]()
    rates = (1.3, 0.7)
    ps = simulate_paths(tree, n, model[0], model[1], seed=5)
    c = _ctx(tree, ps, model)
    R = 24
    try:
        leaf = [v for v in range(tree.n_nodes) if tree.is_leaf(v)]
        end0 = ps.end_state()
        root0 = ps.init_state.reshape(tree.n_nodes, n)[1].copy()
        # analytic E[J], E[D] given the leaves AND the root state, which the sampler keeps
        Je, De = IO.expectation_sufficient_statistics(tree, ps, rates, root=root0)
        Js, Ds = [], []
        for it in range(R):
            c.indep_update_paths(rates, seed=99, iteration=it)
            J, D = c.indep_suffstats()
            Js.append(J); Ds.append(D)
            if it in (0, R - 1):
                out = c.download_paths()
                st = out.init_state.reshape(tree.n_nodes, n)
                assert np.array_equal(out.end_state()[leaf], end0[leaf])       # leaves kept
                assert np.array_equal(st[1], root0)                              # root kept
                e = out.end_state()
                for v in range(1, tree.n_nodes):                                 # rows chain up the tree
                    par = int(tree.parent_ids[v])
                    want = st[1] if par == 0 else e[par]
                    assert np.array_equal(st[v], want)
                off = out.row_off.astype(np.int64)
                for b in range(1, tree.n_nodes):
                    seg = out.jumps[off[b * n]:off[(b + 1) * n]]
                    assert np.all(seg > 0) and np.all(seg < tree.branches[b])
                lens = np.diff(off)
                rows = np.repeat(np.arange(lens.size), lens)
                same = rows[1:] == rows[:-1]
                assert np.all(np.diff(out.jumps)[same] > 0)                      # sorted within a row
        Js, Ds = np.array(Js), np.array(Ds)
        # each iteration is an independent draw given the same leaves: mean over iterations vs expectation
        m = (tree.n_nodes - 1) * 4
        alpha = 1e-3 / m
        zmax = stats.norm.isf(alpha / 2)
        for X, E in ((Js, Je), (Ds, De)):
            mean, se = X.mean(0), X.std(0, ddof=1) / np.sqrt(R)
            z = (mean[1:] - E[1:]) / np.maximum(se[1:], 1e-300)
            assert np.all(np.abs(z) < zmax), (z, zmax)
    finally:
        c.close()


@pytest.mark.skipif(not cpu_libs.ref_available(), reason="oracle/_ref not built")
def test_update_paths_matches_reference_sampler(model):
    tree = ref_test_tree()
    rates = (0.9, 1.6)
    n, R = 3000, 16
    ps = simulate_paths(tree, n, model[0], model[1], seed=15)
    ref = cpu_libs.ref()
    ref.indep_load(ps)
    Jr, Dr = [], []
    for it in range(R):
        ref.indep_update_paths(tree, rates, seed=1000 + it)
        J, D = ref.indep_suffstats()
        Jr.append(J * n); Dr.append(D * n)
    c = _ctx(tree, ps, model)
    Jg, Dg = [], []
    try:
        for it in range(R):
            c.indep_update_paths(rates, seed=7, iteration=it)
            J, D = c.indep_suffstats()
            Jg.append(J); Dg.append(D)
    finally:
        c.close()
    m = (tree.n_nodes - 1) * 4
    for A, B in ((np.array(Jr), np.array(Jg)), (np.array(Dr), np.array(Dg))):
        p = stats.ttest_ind(A[:, 1:].reshape(R, -1), B[:, 1:].reshape(R, -1), equal_var=False).pvalue
        assert np.all(p > 1e-3 / m), p


@pytest.mark.parametrize("world", [2, 3])
def test_update_paths_sharded_equals_single(model, world):
    import torch
    tree = Tree.balanced(8, 0.05)
    rates = (1.3, 0.7)
    n = 1500
    ps = simulate_paths(tree, n, model[0], model[1], seed=4)
    one = _ctx(tree, ps, model)
    one.indep_update_paths(rates, seed=3, iteration=2)
    want = one.download_paths()
    J1, D1 = one.indep_suffstats()
    E1 = one.indep_expectation(rates)
    one.close()
    dev = torch.device("cuda", 0)
    parts, Js, Ds, EJ = [], 0, 0, 0
    for r in range(world):
        c = Context(0)
        try:
            c.set_tree(tree); c.set_model(*model)
            S.Shard(c, n, world, r, dev)
            c.upload_paths(S.local_paths(ps, n, world, r))
            EJ = EJ + c.indep_expectation(rates)[0]
            c.indep_update_paths(rates, seed=3, iteration=2)
            J, D = c.indep_suffstats()
            Js, Ds = Js + J, Ds + D
            parts.append(c.download_paths())
        finally:
            c.close()
    got = S.gather_paths(parts, n, world)
    assert got.same_as(want)
    assert np.array_equal(Js, J1) and np.allclose(Ds, D1, rtol=1e-12)
    assert np.allclose(EJ, E1[0], rtol=1e-12)
