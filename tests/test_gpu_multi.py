"""Several shards behind ONE handle in ONE process (epv_multi_*, csrc/api_multi.cu: the multi-GPU driver
of the C++ drop-in): peer-copied halo records on side streams, edge groups first, overlapped with the
interior update.  Must reproduce the single-context chain bit for bit.  On a one-GPU box the shards
share device 0 (the protocol, streams and events are the same)."""
import os
import subprocess

import numpy as np
import pytest

from epievo_b200.device import Context, MultiContext
from epievo_b200.paths import Tree
from epievo_b200.synthetic import simulate_paths
from synth import ref_test_tree, synth_paths

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def devices(n):
    import torch
    k = torch.cuda.device_count()
    return [r % k for r in range(n)]


@pytest.mark.parametrize("case,n,world", [("testtree", 1000, 2), ("testtree", 1001, 3), ("pair", 500, 4),
                                          ("bal8", 302, 2), ("bal32", 100, 3)])
def test_multi_equals_single(model, case, n, world):
    tree = {"testtree": ref_test_tree, "pair": lambda: Tree.pair(1.0),
            "bal8": lambda: Tree.balanced(8, 0.05), "bal32": lambda: Tree.balanced(32, 0.05)}[case]()
    rates, T = model
    ps = simulate_paths(tree, n, rates, T, seed=3)
    one = Context(0)
    one.set_tree(tree); one.set_model(rates, T); one.upload_paths(ps); one.reset()
    J1, D1, acc1 = one.run_mcmc(2, 3, seed=77, first_sweep=5)
    want = one.download_paths()
    one.close()
    m = MultiContext(world, devices(world))
    try:
        m.set_tree(tree); m.set_model(rates, T); m.upload_paths(ps)
        J, D, acc = m.run_mcmc(2, 3, seed=77, first_sweep=5)
        got = m.download_paths()
        assert got.same_as(want)
        assert np.array_equal(J, J1) and np.allclose(D, D1, rtol=1e-12, atol=0) and acc == acc1
        # J/D of the gathered paths through the multi handle == single context on the same paths
        Js, Ds = m.suffstats()
        chk = Context(0)
        chk.set_tree(tree); chk.upload_paths(want)
        J2, D2 = chk.suffstats(); chk.close()
        assert np.array_equal(Js, J2) and np.allclose(Ds, D2, rtol=1e-12, atol=0)
    finally:
        m.close()


def test_multi_mcem_iteration_with_branch_updates_equals_single(model):
    """The -b path (estimate_rates_and_branches + scale_jump_times every iteration,
    prog/epievo_est_params_histories.cpp:259-264) over 3 shards: same parameters and branch lengths as
    the single context, bit for bit (the M-step sees the same J, and D to the last bits)."""
    tree = ref_test_tree()
    rates, T = model
    ps = simulate_paths(tree, 3000, rates, T, seed=11)
    outs = []
    for world in (1, 3):
        m = MultiContext(world, devices(world))
        try:
            m.set_tree(tree); m.set_model(rates, T); m.upload_paths(ps)
            r, t, br = np.array(rates), np.array(T), np.array(tree.branches, np.float64)
            sweep = 0
            for it in range(3):
                r, t, br, bl, acc, llh = m.mcem_iteration(4, 4, 99, sweep, r, t, br, optimize_branches=True)
                sweep += 8
            outs.append((r, t, br, acc, llh, m.download_paths()))
        finally:
            m.close()
    a, b = outs
    # J and D are folded from integer sums over all shards: the M-step sees identical numbers
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]) and np.array_equal(a[2], b[2])
    assert a[3] == b[3] and a[4] == b[4]
    assert a[5].same_as(b[5])


def test_long_branch_boundary_rows_outgrow_the_default_halo_record(model):
    """T = 20: boundary rows hold ~20-40 jumps and grow past 48 -- the halo-row capacity is sized from
    the data at upload and regrown between sweeps instead of raising ERR_HALO_CAPACITY."""
    tree = Tree.pair(20.0)
    rates, T = model
    ps = synth_paths(tree, 600, 9, 1.0)
    one = Context(0)
    one.set_tree(tree); one.set_model(rates, T); one.upload_paths(ps); one.reset()
    J1, D1, acc1 = one.run_mcmc(3, 2, seed=5)
    want = one.download_paths()
    one.close()
    m = MultiContext(2, devices(2))
    try:
        m.set_tree(tree); m.set_model(rates, T); m.upload_paths(ps)
        J, D, acc = m.run_mcmc(3, 2, seed=5)
        assert m.download_paths().same_as(want)
        assert np.array_equal(J, J1) and acc == acc1
    finally:
        m.close()


def test_cpp_host_program_drives_two_contexts():
    """tests/cpp/test_multi.cpp: plain C++ over the C-ABI, 2 and 3 shards vs 1."""
    exe = os.path.join(ROOT, "tests", "cpp", "_build", "test_multi")
    if not os.path.exists(exe):
        pytest.skip("tests/cpp/_build/test_multi not built (__graft_entry__.build())")
    for shards in ("2", "3"):
        r = subprocess.run([exe, shards, "3000"], capture_output=True, text=True)
        assert r.returncode == 0, r.stdout + r.stderr
        assert r.stdout.startswith("ok"), r.stdout
