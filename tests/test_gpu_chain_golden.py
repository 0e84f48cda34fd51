"""Regression pin of whole chains: the histories after run_mcmc(3, 3) on fixed inputs must equal, bit for
bit, the ones recorded with the round-1 kernels (tests/golden/chain_golden.npz, written by
tools/record_chain_golden.py on a B200): same Philox streams, same arithmetic.  J exact; D and the
cached tri_llh to 1e-12 (the update kernel may sum the same terms in another grouping)."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
GOLD = os.path.join(ROOT, "tests", "golden", "chain_golden.npz")

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not os.path.exists(GOLD), reason="no recorded chains")]

import record_chain_golden as rec


@pytest.mark.parametrize("case", rec.CASES, ids=[c[0] for c in rec.CASES])
def test_chain_equals_recorded_chain(ctx, case):
    gold = np.load(GOLD)
    name = case[0]
    got = rec.run_case(ctx, *rec.make_case(*case))
    assert int(got["n_jumps"]) == int(gold[name + ".n_jumps"])
    assert np.array_equal(got["sha"], gold[name + ".sha"]), "histories differ from the recorded chain"
    assert float(got["acc"]) == float(gold[name + ".acc"])
    assert np.array_equal(got["J"], gold[name + ".J"])
    assert np.allclose(got["D"], gold[name + ".D"], rtol=1e-12, atol=0)
    assert np.allclose(got["tri"], gold[name + ".tri"], rtol=1e-12, atol=1e-13)
