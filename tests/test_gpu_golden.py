"""CUDA path vs the committed golden vectors of the reference build (tests/golden/):
same inputs, integer counts exact, floating point within 1e-9 relative."""
import numpy as np
import pytest

from synth import rel_close
from test_golden_oracle import CASES, SITES, load_case

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", CASES)
def test_cuda_matches_reference_golden(ctx, name):
    g, tree, ps = load_case(name)
    rates, T = g["rates"], g["T"]
    ctx.set_tree(tree); ctx.set_model(rates, T); ctx.upload_paths(ps)
    J, D = ctx.suffstats()
    assert np.array_equal(J, g["J"])
    assert rel_close(D, g["D"], 1e-9)
    ctx.reset()
    assert rel_close(ctx.tri_llh(), g["tri_llh"], 1e-9, atol=1e-12)
    for s in SITES:
        d = ctx.site_debug(s)
        for k in ("seg_off", "trip0", "trip1"):
            assert np.array_equal(d[k], g["dbg%d_%s" % (s, k)])
        for k in ("rate0", "rate1", "len", "p", "q", "root_p0"):
            assert rel_close(d[k], g["dbg%d_%s" % (s, k)], 1e-9, atol=1e-300), (s, k)
        want = float(g["dbg%d_proposal_logp" % s])
        assert abs(d["proposal_logp"] - want) <= 1e-9*max(1.0, abs(want))
    ctx.scale_jump_times(g["scale_branches"])
    assert rel_close(ctx.download_paths().jumps, g["scaled_jumps"], 1e-15)


@pytest.mark.parametrize("name", CASES)
def test_cuda_stats_on_reference_chain_output(ctx, name):
    """J/D of the paths the reference's own sampler produced (golden chain end state)."""
    from epievo_b200.paths import PathSet
    import cpu_libs
    g, tree, ps = load_case(name)
    out = PathSet(ps.n_sites, ps.n_nodes, g["mcmc_init"], g["mcmc_off"], g["mcmc_jumps"], ps.tot_time)
    ctx.set_tree(tree); ctx.set_model(g["rates"], g["T"]); ctx.upload_paths(out)
    J, D = ctx.suffstats()
    o = cpu_libs.oracle(); o.load_paths(out)
    Jr, Dr = o.get_suff_stats()
    assert np.array_equal(J, Jr) and rel_close(D, Dr, 1e-9)
