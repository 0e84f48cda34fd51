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
