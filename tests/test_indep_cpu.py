"""Site-independent E-step of epievo_initialization, CPU side:
* the numpy restatement (oracle/indep_oracle.py) is pinned against the UNMODIFIED reference
  (IndepSite.cpp via oracle/_ref): expectation_sufficient_statistics and
  compute_sufficient_statistics on 4 trees;
* the library's host functions (tables, estimate_rates_indep, ..._and_branches_indep) against
  the reference's."""
import ctypes as C
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "oracle"))
import cpu_libs  # noqa: E402
import indep_oracle as IO  # noqa: E402
from epievo_b200 import _lib as L  # noqa: E402
from epievo_b200.paths import Tree  # noqa: E402
from epievo_b200.synthetic import simulate_paths  # noqa: E402
from synth import ref_test_tree  # noqa: E402

TREES = {"testtree": ref_test_tree, "pair": lambda: Tree.pair(1.0),
         "bal8": lambda: Tree.balanced(8, 0.05), "bal32": lambda: Tree.balanced(32, 0.3)}
RATES = [(1.3, 0.7), (0.2, 2.5)]

needs_ref = pytest.mark.skipif(not cpu_libs.ref_available(), reason="oracle/_ref not built")


@needs_ref
@pytest.mark.parametrize("case", sorted(TREES))
@pytest.mark.parametrize("rates", RATES)
def test_numpy_restatement_equals_reference(model, case, rates):
    tree = TREES[case]()
    ps = simulate_paths(tree, 400, model[0], model[1], seed=21)
    ref = cpu_libs.ref()
    ref.indep_load(ps)
    Jr, Dr = ref.indep_expectation(tree, rates)
    J, D = IO.expectation_sufficient_statistics(tree, ps, rates)
    assert np.allclose(J, Jr, rtol=1e-12, atol=1e-13) and np.allclose(D, Dr, rtol=1e-12, atol=1e-13)
    Jr, Dr = ref.indep_suffstats()                       # averages
    J, D = IO.compute_sufficient_statistics(ps)          # sums
    assert np.array_equal(J / ps.n_sites, Jr)
    assert np.allclose(D / ps.n_sites, Dr, rtol=1e-12, atol=0)


def _d(a):
    return a.ctypes.data_as(L.dp)


@pytest.mark.parametrize("r0,r1,T", [(1.5, 0.5, 1.0), (0.3, 2.0, 0.05), (4.0, 4.0, 0.7)])
def test_host_tables(r0, r1, T):
    lib = L.load()
    out = [np.zeros(4) for _ in range(5)]
    lib.epv_indep_tables(r0, r1, T, *[_d(x) for x in out])
    P, J0, J1, D0, D1 = [x.reshape(2, 2) for x in out]
    assert np.array_equal(P, IO.trans_prob_mat(r0, r1, T))
    a, b = IO.expectation_J(r0, r1, T)
    assert np.allclose(J0, a, rtol=1e-14) and np.allclose(J1, b, rtol=1e-14)
    a, b = IO.expectation_D(r0, r1, T)
    assert np.allclose(D0, a, rtol=1e-14) and np.allclose(D1, b, rtol=1e-14)
    # known answers printed by the reference's end_cond_sampling_test
    if (r0, r1, T) == (1.5, 0.5, 1.0):
        assert abs(J0[0, 0] - 0.389) < 1e-3 and abs(D0[0, 0] - 0.856) < 1e-3
        assert abs(J0[0, 1] - 1.12) < 1e-2 and abs(J1[0, 1] - 0.117) < 1e-3 and abs(D0[0, 1] - 0.422) < 1e-3


@needs_ref
@pytest.mark.parametrize("with_branches", [False, True])
def test_host_m_step_equals_reference(model, with_branches):
    tree = ref_test_tree()
    ps = simulate_paths(tree, 300, model[0], model[1], seed=3)
    ref = cpu_libs.ref()
    ref.indep_load(ps)
    J, D = ref.indep_suffstats()
    want_r, want_br = ref.indep_estimate(tree, J, D, (0.0, 0.0), with_branches)
    lib = L.load()
    r = np.zeros(2)
    br = np.ascontiguousarray(tree.branches, np.float64).copy()
    Jf, Df = np.ascontiguousarray(J.reshape(-1)), np.ascontiguousarray(D.reshape(-1))
    if with_branches:
        lib.epv_indep_estimate_rates_and_branches(tree.n_nodes, _d(Jf), _d(Df), _d(r), _d(br))
    else:
        lib.epv_indep_estimate_rates(tree.n_nodes, _d(Jf), _d(Df), _d(r))
    assert np.allclose(r, want_r, rtol=1e-14) and np.allclose(br, want_br, rtol=1e-14)
    if with_branches:                      # the reference also rescaled its paths
        got = ref.indep_store()
        assert np.allclose(got.tot_time[1:], br[1:], rtol=1e-14)
