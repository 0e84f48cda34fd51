"""Deterministic parity of the CUDA path against the CPU checker on identical paths
(integer counts exact, floating point within 1e-9 relative)."""
import numpy as np
import pytest

from epievo_b200.paths import Tree
from synth import rel_close, synth_paths, ref_test_tree

pytestmark = pytest.mark.gpu

CASES = [("testtree", 3000, 1.0), ("testtree_dense", 500, 25.0), ("pair", 2000, 1.0),
         ("bal8", 700, 2.0), ("bal32", 200, 3.0),
         # sizes around the 128-row tiles of suffstats_tile_kernel (first tile starts at row 4)
         ("testtree", 133, 2.0), ("testtree", 134, 2.0), ("testtree", 135, 8.0), ("testtree", 262, 1.0),
         ("bal8", 1029, 1.0), ("bal32", 5000, 1.0), ("testtree", 40000, 1.0)]


def make(case, n, density, seed=11):
    tree = {"testtree": ref_test_tree, "testtree_dense": ref_test_tree,
            "pair": lambda: Tree.pair(1.0), "bal8": lambda: Tree.balanced(8, 0.05),
            "bal32": lambda: Tree.balanced(32, 0.05)}[case]()
    return tree, synth_paths(tree, n, seed, density)


@pytest.mark.parametrize("case,n,density", CASES)
def test_roundtrip_and_suffstats(ctx, cpu, model, case, n, density):
    tree, ps = make(case, n, density)
    ctx.set_tree(tree); ctx.set_model(*model); ctx.upload_paths(ps)
    back = ctx.download_paths()
    assert back.same_as(ps)
    cpu.load_paths(ps)
    Jr, Dr = cpu.get_suff_stats()
    J, D = ctx.suffstats()
    assert np.array_equal(J, Jr)                      # integer counts: exact
    assert rel_close(D, Dr, 1e-9)
    llh, g = ctx.llh_grad(model[0])
    import cpu_libs
    o = cpu_libs.oracle()
    Jc, Dc = Jr[1:].sum(0), Dr[1:].sum(0)
    assert rel_close(llh, o.log_likelihood(Jc, Dc, model[0]), 1e-9)
    assert rel_close(g, o.gradient(Jc, Dc, model[0]), 1e-9, atol=1e-9*np.abs(Dc).max())


@pytest.mark.parametrize("case,n,density", CASES)
def test_tri_llh_and_site_intermediates(ctx, cpu, model, case, n, density):
    tree, ps = make(case, n, density)
    rates, T = model
    ctx.set_tree(tree); ctx.set_model(rates, T); ctx.upload_paths(ps); ctx.reset()
    cpu.load_paths(ps)
    assert rel_close(ctx.tri_llh(), cpu.tri_llh(rates, T), 1e-9, atol=1e-12)
    for site in [1, 2, n // 3, n // 2, n - 2]:
        a = ctx.site_debug(site); b = cpu.site_debug(tree, rates, T, site)
        assert np.array_equal(a["seg_off"], b["seg_off"])
        assert np.array_equal(a["trip0"], b["trip0"]) and np.array_equal(a["trip1"], b["trip1"])
        assert np.array_equal(a["len"], b["len"])               # same subtraction: bit-exact
        assert np.array_equal(a["rate0"], b["rate0"]) and np.array_equal(a["rate1"], b["rate1"])
        assert rel_close(a["p"], b["p"], 1e-9) and rel_close(a["q"], b["q"], 1e-9)
        assert rel_close(a["root_p0"], b["root_p0"], 1e-9)
        assert abs(a["proposal_logp"] - b["proposal_logp"]) <= 1e-9*max(1.0, abs(b["proposal_logp"]))


def test_scale_jump_times(ctx, cpu, model):
    tree, ps = make("testtree", 1500, 5.0)
    ctx.set_tree(tree); ctx.set_model(*model); ctx.upload_paths(ps)
    nb = np.array(tree.branches) * np.array([1.0, 1.3, 0.7, 2.0, 0.9])
    ctx.scale_jump_times(nb)
    cpu.load_paths(ps); cpu.scale_jump_times(nb)
    assert ctx.download_paths().same_as(cpu.store_paths())       # one multiply: bit-exact


@pytest.mark.parametrize("case,n,levels,p", [("testtree", 1500, 6, 0.25), ("testtree", 700, 3, 0.6),
                                             ("pair", 2000, 4, 0.5), ("bal8", 600, 5, 0.15)])
def test_suffstats_with_tied_jump_times(ctx, cpu, model, case, n, levels, p):
    """Jump times on a coarse grid: neighbouring rows hold exactly equal times all the time, which
    pins the tie rule of the J/D kernels (the higher site jumps first, Path.cpp:275-281)."""
    from test_jd_identity_cpu import grid_paths
    tree = {"testtree": ref_test_tree, "pair": lambda: Tree.pair(1.0),
            "bal8": lambda: Tree.balanced(8, 0.05)}[case]()
    ps = grid_paths(tree, n, 23, levels, p)
    ctx.set_tree(tree); ctx.set_model(*model); ctx.upload_paths(ps)
    cpu.load_paths(ps)
    Jr, Dr = cpu.get_suff_stats()
    J, D = ctx.suffstats()
    assert np.array_equal(J, Jr)
    assert rel_close(D, Dr, 1e-9)
    ctx.reset()
    assert rel_close(ctx.tri_llh(), cpu.tri_llh(*model), 1e-9, atol=1e-12)
