"""The GPU sampler against the EXACT posterior (the reference harness MCMC_test's design,
tests/exact_posterior.py): 2e4 forward-simulation + rejection draws from the unmodified reference
(oracle/_ref/bin/MCMC_test) vs 2e4 draws of the colour-class sampler, 10 sweeps apart, chi-square on
every J_t histogram and KS on every D_t, family-wise alpha = 1e-3 per case.  The reference's own
CPU sampler goes through the same test (tests/test_exact_posterior_cpu.py) so that a failure here
can be attributed."""
import os

import numpy as np
import pytest

import exact_posterior as ep

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not os.path.exists(ep.REF_MCMC_TEST), reason="oracle/_ref not built")]

N_DRAWS, THIN = 20000, 10


def gpu_chain(ctx, model, root, leaf, T, n_draws, thin, seed):
    rates, Tm = model
    tree, ps = ep.initial_paths(root, leaf, T, seed)
    ctx.set_tree(tree); ctx.set_model(rates, Tm); ctx.upload_paths(ps); ctx.reset()
    out = np.zeros((n_draws, 16))
    sweep = 0
    for k in range(n_draws):
        J, D, _ = ctx.run_mcmc(thin, 1, seed=seed, first_sweep=sweep)      # MCMC_test.cpp:394-403
        sweep += thin + 1
        out[k, :8] = J[1]; out[k, 8:] = D[1]
    return out


@pytest.mark.parametrize("n,T,seed", [(5, 0.1, 7), (5, 1.0, 7), (8, 1.0, 7), (12, 0.1, 3), (12, 1.0, 5)])
def test_gpu_sampler_matches_exact_posterior(ctx, model, tmp_path, n, T, seed):
    root, leaf, exact, ref_chain = ep.run_harness(ep.REF_MCMC_TEST, n, T, N_DRAWS, THIN, seed, str(tmp_path / "ref"))
    chain = gpu_chain(ctx, model, root, leaf, T, N_DRAWS, THIN, 900 + seed)
    # leaf data and the fixed end sites survive 2.2e5 sweeps
    out = ctx.download_paths()
    assert np.array_equal(out.end_state()[1], leaf) and np.array_equal(out.init_state[n:], root)
    p_ref = ep.compare(exact, ref_chain)
    p = ep.assert_same_law(exact, chain, "GPU sampler n=%d T=%g (reference sampler's smallest p: %.3g)"
                           % (n, T, min(p_ref.values())))
    print("n=%d T=%g smallest p: gpu %.3g, reference %.3g" % (n, T, min(p.values()), min(p_ref.values())))


def test_dropin_harness_matches_exact_posterior(tmp_path):
    """The reference's harness itself, compiled unchanged against the drop-in header and linked with
    libepievo_b200.so (integration/Makefile): its .mcmc file against its own .forward file."""
    if not os.path.exists(ep.B200_MCMC_TEST):
        pytest.skip("integration/_build not built")
    _, _, exact, chain = ep.run_harness(ep.B200_MCMC_TEST, 8, 1.0, 4000, 10, 7, str(tmp_path / "b200"))
    ep.assert_same_law(exact, chain, "MCMC_test_b200 n=8 T=1")


# ---- trees: exact draws from tests/exact_tree.py (pinned to the reference harness on one branch) ------
import exact_tree as et
from test_exact_posterior_cpu import compare_tree, cpu_chain, tree_case


@pytest.mark.parametrize("name,n,seed", [("testtree_x5", 6, 6), ("testtree_x5", 6, 4), ("cherry", 7, 4)])
def test_gpu_sampler_matches_exact_posterior_on_a_tree(ctx, cpu, model, name, n, seed):
    """Internal nodes, several children, per-branch statistics: 2e4 exact whole-tree draws (forward
    simulation along every branch + rejection on all leaf sequences) vs 2e4 draws of the GPU sampler,
    chi-square / KS on J_t / D_t of every branch, family-wise alpha = 1e-3; the reference's CPU sampler
    through the same test for attribution."""
    tree, root, leaves = tree_case(name, n, model, seed)
    rates, T = model
    exact, acc = et.exact_draws(tree, root, leaves, rates, N_DRAWS, seed=100 + seed)
    ps = et.initial_paths(tree, root, leaves, seed)
    ctx.set_tree(tree); ctx.set_model(rates, T); ctx.upload_paths(ps); ctx.reset()
    chain = np.zeros((N_DRAWS, tree.n_nodes, 16))
    sweep = 0
    for k in range(N_DRAWS):
        J, D, _ = ctx.run_mcmc(THIN, 1, seed=4242 + seed, first_sweep=sweep)
        sweep += THIN + 1
        chain[k, :, :8] = J; chain[k, :, 8:] = D
    p_ref = compare_tree(exact, cpu_chain(cpu, tree, model, ps, N_DRAWS, THIN, 77), "reference sampler on " + name)
    p = compare_tree(exact, chain, "GPU sampler on %s (reference sampler's smallest p: %.3g)" % (name, p_ref))
    print("%s seed %d: exact-sampler acceptance %.3g; smallest p: gpu %.3g, reference %.3g" % (name, seed, acc, p, p_ref))
