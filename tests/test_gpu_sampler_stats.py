"""Statistical parity of the stride-3 colour-class GPU sampler with the reference's
sequential sampler (SingleSiteSampler::run_mcmc): both target the same posterior over
histories given the leaf states and the two fixed boundary sites, so independent chains
from each must give the same distribution of every summary statistic.

Stated tests (all two-sided, R independent chains per arm, same starting paths):
  * per (branch, triplet) batch-mean J and D:   two-sample Kolmogorov-Smirnov,
    and Welch's t-test, each Bonferroni-corrected over all entries at family-wise
    alpha = 1e-3 (tools/stats_probe.py is the same comparison at R = 96: max |z| = 3.0
    over 224 entries, mean z^2 = 1.24);
  * per-branch site marginals (fraction of sites in state 1 at 1/4, 1/2, 3/4 of each
    branch): the same KS / z tests;
  * site-resolved marginals at mid-branch: Welch z per (branch, site) over the sites whose
    pooled frequency is in (0.05, 0.95), chain-level variance: mean z^2 < 1.6 (null
    expectation nu/(nu-2) ~ 1.07) and max z^2 < 60.
"""
import numpy as np
import pytest
from scipy import stats

from epievo_b200.paths import Tree
from epievo_b200.synthetic import simulate_paths
from synth import ref_test_tree

pytestmark = pytest.mark.gpu

R = 16          # chains per arm
BURN = 30
DRAWS = 8       # path draws per chain, 3 sweeps apart


def states_at(ps, frac):
    """[n_nodes, n_sites] state of every row at time frac*tot_time."""
    lens = ps.row_len().reshape(-1)
    rid = np.repeat(np.arange(lens.size), lens)
    t = np.repeat(ps.tot_time, ps.n_sites)*frac
    cnt = np.bincount(rid[ps.jumps < t[rid]], minlength=lens.size)
    return ((ps.init_state + cnt) & 1).reshape(ps.n_nodes, ps.n_sites)


def summarise(draws):
    """draws: list of PathSet -> (branch-level marginals [B*3], per-site mid marginals [B, n])"""
    agg = np.mean([[states_at(p, f)[1:, 1:-1].mean(1) for f in (0.25, 0.5, 0.75)] for p in draws], 0)
    mid = np.mean([states_at(p, 0.5)[1:] for p in draws], 0)
    return agg.reshape(-1), mid


def compare(a, b, label):
    """a, b: [R, k] per-chain statistics."""
    k = a.shape[1]
    alpha = 1e-3/k
    for i in range(k):
        if a[:, i].std() == 0 and b[:, i].std() == 0:
            assert a[0, i] == b[0, i], (label, i)
            continue
        p = stats.ks_2samp(a[:, i], b[:, i]).pvalue
        pt = stats.ttest_ind(a[:, i], b[:, i], equal_var=False).pvalue
        assert p > alpha, "%s[%d]: KS p=%.2e (ref %.5g vs gpu %.5g)" % (label, i, p, a[:, i].mean(), b[:, i].mean())
        assert pt > alpha, "%s[%d]: Welch p=%.2e (ref %.5g vs gpu %.5g)" % (label, i, pt, a[:, i].mean(), b[:, i].mean())


@pytest.mark.parametrize("case,n", [("testtree", 2000), ("pair", 600), ("bal8", 600)])
def test_gpu_chain_matches_reference_chain(ctx, cpu, model, case, n):
    tree = {"testtree": ref_test_tree, "pair": lambda: Tree.pair(1.0),
            "bal8": lambda: Tree.balanced(8, 0.05)}[case]()
    rates, T = model
    ps = simulate_paths(tree, n, rates, T, seed=77)
    ctx.set_tree(tree); ctx.set_model(rates, T)
    ref_JD, gpu_JD, ref_agg, gpu_agg, ref_mid, gpu_mid = [], [], [], [], [], []
    for c in range(R):
        # reference arm: sequential sweeps driven by mt19937(seed)
        cpu.load_paths(ps); cpu.seed(1000 + c)
        J, D, _, _ = cpu.run_mcmc(tree, rates, T, BURN, 10)
        ref_JD.append(np.concatenate([J[1:].ravel(), D[1:].ravel()]))
        draws = []
        for _ in range(DRAWS):
            cpu.run_mcmc(tree, rates, T, 2, 1, reset=False)
            draws.append(cpu.store_paths())
        a, m = summarise(draws); ref_agg.append(a); ref_mid.append(m)
        # GPU arm: colour-class sweeps driven by Philox(seed)
        ctx.upload_paths(ps); ctx.reset()
        J, D, _ = ctx.run_mcmc(BURN, 10, seed=5000 + c)
        gpu_JD.append(np.concatenate([J[1:].ravel(), D[1:].ravel()]))
        draws, sw = [], BURN + 10
        for _ in range(DRAWS):
            for _ in range(3):
                ctx.sweep(5000 + c, sw); sw += 1
            draws.append(ctx.download_paths())
        a, m = summarise(draws); gpu_agg.append(a); gpu_mid.append(m)
    compare(np.array(ref_JD), np.array(gpu_JD), case + " J/D")
    compare(np.array(ref_agg), np.array(gpu_agg), case + " marginals")
    # site-resolved check at mid-branch: Welch z per (branch, site) with the variance
    # estimated ACROSS chains (draws within a chain are autocorrelated), mean z^2 ~ 1
    rm, gm = np.array(ref_mid), np.array(gpu_mid)          # [R, B, n]
    pr, pg = rm.mean(0), gm.mean(0)
    pooled = 0.5*(pr + pg)
    sel = (pooled > 0.05) & (pooled < 0.95)
    sel[:, 0] = sel[:, -1] = False
    m = int(sel.sum())
    if m >= 20:
        se2 = rm.var(0, ddof=1)/R + gm.var(0, ddof=1)/R
        z2 = ((pr - pg)**2)[sel]/se2[sel]
        assert z2.mean() < 1.6, "%s: site-resolved marginals mean z^2 = %.2f over %d cells" % (case, z2.mean(), m)
        assert z2.max() < 60.0, "%s: site-resolved marginals max z^2 = %.1f" % (case, z2.max())
