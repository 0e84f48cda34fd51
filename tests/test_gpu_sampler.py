"""Structural checks of the Metropolis-Hastings sweep on the GPU: every invariant the
reference's update preserves (SingleSiteSampler.cpp:482-536) must hold after sweeps,
and the cached tri_llh must equal a from-scratch recomputation."""
import numpy as np
import pytest

from epievo_b200.paths import Tree
from synth import rel_close, synth_paths, ref_test_tree

pytestmark = pytest.mark.gpu


def check_valid(tree, before, after):
    n = before.n_sites
    assert (after.n_sites, after.n_nodes) == (n, before.n_nodes)
    st0 = before.init_state.reshape(before.n_nodes, n)
    st1 = after.init_state.reshape(after.n_nodes, n)
    e0, e1 = before.end_state(), after.end_state()
    for v in range(1, tree.n_nodes):
        par = int(tree.parent_ids[v])
        if par == 0:
            assert np.array_equal(st1[v], st0[1])          # root state never resampled (:441)
        else:
            assert np.array_equal(st1[v][1:-1], e1[par][1:-1])   # branch starts at parent's end
        if tree.is_leaf(v):
            assert np.array_equal(e1[v], e0[v])            # leaf data preserved
        T = float(tree.branches[v])
        for s in range(n):
            r = after.row(v, s)
            if len(r):
                assert r[0] > 0.0 and r[-1] < T and np.all(np.diff(r) > 0)
    # boundary sites are never updated (:545)
    for v in range(1, tree.n_nodes):
        assert np.array_equal(after.row(v, 0), before.row(v, 0))
        assert np.array_equal(after.row(v, n - 1), before.row(v, n - 1))


@pytest.mark.parametrize("case", ["testtree", "pair", "bal8", "bal32"])
def test_sweeps_preserve_invariants(ctx, cpu, model, case):
    tree = {"testtree": ref_test_tree, "pair": lambda: Tree.pair(1.0),
            "bal8": lambda: Tree.balanced(8, 0.05), "bal32": lambda: Tree.balanced(32, 0.05)}[case]()
    n = 1000
    ps = synth_paths(tree, n, 5, 1.0)
    rates, T = model
    ctx.set_tree(tree); ctx.set_model(rates, T); ctx.upload_paths(ps); ctx.reset()
    acc = 0
    for sw in range(4):
        acc += ctx.sweep(1234, sw)
    assert 0.5*4*(n - 2) < acc <= 4*(n - 2)
    out = ctx.download_paths()
    check_valid(tree, ps, out)
    assert not out.same_as(ps)
    cached = ctx.tri_llh()
    cpu.load_paths(out)
    assert rel_close(cached, cpu.tri_llh(rates, T), 1e-9, atol=1e-12)
    # determinism: same seed, same result
    ctx.upload_paths(ps); ctx.reset()
    for sw in range(4):
        ctx.sweep(1234, sw)
    assert ctx.download_paths().same_as(out)


def test_end_cond_sampler_matches_analytic_expectations(ctx, cpu):
    """Known answers: expectation_J / expectation_D
    (ContinuousTimeMarkovModel.cpp:167-226), z-test at 5 sigma on 4e5 draws."""
    r0, r1, T, n = 1.5, 0.5, 1.0, 400000
    J0, J1, D0, D1 = cpu.expectations(r0, r1, T)
    for a in (0, 1):
        for b in (0, 1):
            nj, j = ctx.end_cond_sample(r0, r1, a, b, T, 99 + 2*a + b, n, cap=48)
            assert nj.max() < 48 and np.all((nj & 1) == (a != b))
            k = np.arange(48)[None, :]
            valid = k < nj[:, None]
            # time in state a = sum of even segments; count jumps out of 0 / out of 1
            t = np.where(valid, j, T)
            seg = np.diff(np.concatenate([np.zeros((n, 1)), t, np.full((n, 1), T)], axis=1), axis=1)
            even = (np.arange(seg.shape[1]) % 2 == 0)[None, :]
            time_a = (seg*even).sum(1)
            d0 = time_a if a == 0 else T - time_a
            from_a = (nj + 1)//2
            from_b = nj//2
            j0 = from_a if a == 0 else from_b
            j1 = from_b if a == 0 else from_a
            for got, want in ((d0, D0[2*a + b]), (j0, J0[2*a + b]), (j1, J1[2*a + b])):
                se = got.std()/np.sqrt(n) + 1e-12
                assert abs(got.mean() - want) < 5*se, (a, b, got.mean(), want, se)


def _end_cond_moments(nj, j, a, T, cap):
    """Per draw: time spent in state 0, jumps out of 0, jumps out of 1."""
    n = nj.shape[0]
    k = np.arange(cap)[None, :]
    t = np.where(k < nj[:, None], j, T)
    seg = np.diff(np.concatenate([np.zeros((n, 1)), t, np.full((n, 1), T)], axis=1), axis=1)
    even = (np.arange(seg.shape[1]) % 2 == 0)[None, :]
    time_a = (seg*even).sum(1)
    d0 = time_a if a == 0 else T - time_a
    from_a, from_b = (nj + 1)//2, nj//2
    return d0, (from_a if a == 0 else from_b), (from_b if a == 0 else from_a)


@pytest.mark.parametrize("r0,r1", [(0.09, 10.2), (10.2, 0.09), (0.5, 0.5), (3.0, 0.3), (10.2, 10.2), (0.09, 0.09)])
@pytest.mark.parametrize("T", [0.01, 0.05, 0.1, 1.0])
def test_end_cond_sampler_parameter_sweep(ctx, cpu, r0, r1, T):
    """The same known answers over the context rates the model actually produces (0.09-10.2 with
    test.param) and over production interval lengths (0.01-0.1) as well as T = 1: E[J0], E[J1], E[D0]
    given both end states (expectation_J / expectation_D, ContinuousTimeMarkovModel.cpp:167-226).
    z-test at 5 sigma per statistic (288 statistics over the sweep: family-wise 1.7e-4)."""
    n, cap = 200000, 64
    J0, J1, D0, D1 = cpu.expectations(r0, r1, T)
    for a in (0, 1):
        for b in (0, 1):
            nj, j = ctx.end_cond_sample(r0, r1, a, b, T, 1234 + 2*a + b, n, cap=cap)
            assert nj.max() < cap and np.all((nj & 1) == (a != b))
            jj = np.where(np.arange(cap)[None, :] < nj[:, None], j, np.inf)
            assert np.all(np.diff(np.concatenate([np.zeros((n, 1)), jj], 1), axis=1)[np.isfinite(jj)] > 0)
            assert np.all(j[np.arange(cap)[None, :] < nj[:, None]] < T)
            d0, j0, j1 = _end_cond_moments(nj, j, a, T, cap)
            for name, got, want in (("D0", d0, D0[2*a + b]), ("J0", j0, J0[2*a + b]), ("J1", j1, J1[2*a + b])):
                se = max(got.std()/np.sqrt(n), 1.0/n)     # (rare events: never below one draw's worth)
                assert abs(got.mean() - want) < 5*se + 1e-12, (r0, r1, T, a, b, name, got.mean(), want, se)


@pytest.mark.parametrize("case,n,density", [("testtree", 3000, 1.0), ("testtree", 600, 25.0), ("bal8", 1500, 6.0),
                                            ("pair", 2000, 8.0)])
def test_deferred_overflow_launch_gives_the_same_chain(model, case, n, density, monkeypatch):
    """Long passes defer the sites the flat routine cannot finish (proposal or interval stack too
    large) to a second launch of the full routine (kernels_mh.cu, mh_update_group); a site's result
    depends on its data only, so the chain must be bit-identical either way.  The dense cases make
    many sites overflow."""
    from epievo_b200.device import Context
    tree = {"testtree": ref_test_tree, "pair": lambda: Tree.pair(1.0),
            "bal8": lambda: Tree.balanced(8, 0.05)}[case]()
    ps = synth_paths(tree, n, 5, density)
    out = []
    for defer_min in ("0", "1000000000000"):
        monkeypatch.setenv("EPIEVO_B200_DEFER_MIN", defer_min)
        c = Context(0)
        try:
            c.set_tree(tree); c.set_model(*model); c.upload_paths(ps); c.reset()
            J, D, acc = c.run_mcmc(2, 2, seed=9)
            out.append((J, D, acc, c.download_paths(), c.tri_llh()))
        finally:
            c.close()
    a, b = out
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]) and a[2] == b[2]
    assert a[3].same_as(b[3]) and np.array_equal(a[4], b[4])


@pytest.mark.parametrize("case,T", [("pair", 3.0), ("pair", 20.0), ("testtree_x10", None)])
def test_growth_beyond_the_jump_capacity_is_recovered(cpu, model, case, T):
    """Histories that start with 0/1 jumps per row on long branches (the heuristic starting paths of
    epievo_initialization / MCMC_test.cpp:219-237) grow several-fold within the first sweeps: a colour pass
    then outgrows the jump storage sized at upload.  The rebuild's inputs are intact, so the library regrows
    and reruns it (epv_recovery_count) instead of failing with "jump storage capacity exceeded".
    EPIEVO_B200_TIGHT_CAPACITY (test-only) sizes the storage at upload with 3 % headroom so that the growth
    certainly goes through that path; the chain must equal the one of a run with the normal headroom."""
    from epievo_b200.device import Context
    from epievo_b200.paths import PathSet, parse_newick
    rates, Tm = model
    tree = Tree.pair(T) if case == "pair" else parse_newick("((C:0.3,D:0.6)E:0.2,F:1.0)G:0.0;")
    n = 30000
    rng = np.random.default_rng(2)
    # every node's sequence: a coarse random pattern; a branch gets one jump where its two ends differ
    seq = [(rng.random(n) < 0.5).astype(np.uint8) for _ in range(tree.n_nodes)]
    rows, init = [[[] for _ in range(n)]], [np.zeros(n, np.uint8)]
    for v in range(1, tree.n_nodes):
        a = seq[int(tree.parent_ids[v])]
        u = rng.random(n)*float(tree.branches[v])
        rows.append([[u[s]] if a[s] != seq[v][s] else [] for s in range(n)])
        init.append(a)
    tot = np.array(tree.branches, np.float64); tot[0] = 0.0
    ps = PathSet.from_rows(n, tree.n_nodes, init, rows, tot)
    outs = []
    for rep in range(2):
        c = Context(0)
        try:
            c.set_tree(tree); c.set_model(rates, Tm); c.upload_paths(ps); c.reset()
            J, D, acc = c.run_mcmc(6, 2, seed=4)
            out = c.download_paths()
            assert out.jumps.size > 1.5*ps.jumps.size
            cached = c.tri_llh()
            outs.append((J, D, acc, out))
        finally:
            c.close()
    check_valid(tree, ps, outs[0][3])
    cpu.load_paths(outs[0][3])
    assert rel_close(cached, cpu.tri_llh(rates, Tm), 1e-9, atol=1e-12)
    # time is conserved, and the run is reproducible
    assert np.allclose(outs[0][1][1:].sum(1), (n - 2)*np.asarray(tree.branches)[1:], rtol=1e-9)
    assert outs[0][3].same_as(outs[1][3]) and np.array_equal(outs[0][0], outs[1][0])
    # the same problem in a fresh process with tight storage: recovered, and the same chain
    import json, subprocess, sys, tempfile, os
    with tempfile.TemporaryDirectory() as tmp:
        np.savez(os.path.join(tmp, "in.npz"), init=ps.init_state, off=ps.row_off, jumps=ps.jumps, tot=ps.tot_time,
                 ss=tree.subtree_sizes, par=tree.parent_ids, br=tree.branches, rates=rates, T=Tm)
        code = (
            "import sys, json, numpy as np\n"
            "sys.path.insert(0, %r)\n"
            "from epievo_b200.device import Context\n"
            "from epievo_b200.paths import PathSet, Tree\n"
            "d = np.load(%r)\n"
            "tree = Tree(d['ss'], d['par'], d['br'], ['n%%d' %% i for i in range(len(d['br']))])\n"
            "n = len(d['init'])//tree.n_nodes\n"
            "ps = PathSet(n, tree.n_nodes, d['init'], d['off'], d['jumps'], d['tot'])\n"
            "c = Context(0); c.set_tree(tree); c.set_model(d['rates'], d['T']); c.upload_paths(ps); c.reset()\n"
            "J, D, acc = c.run_mcmc(6, 2, seed=4)\n"
            "out = c.download_paths()\n"
            "np.savez(%r, jumps=out.jumps, off=out.row_off, init=out.init_state, J=J)\n"
            "print(json.dumps({'recoveries': c.recovery_count()}))\n"
        ) % (os.path.dirname(os.path.dirname(os.path.abspath(__file__))), os.path.join(tmp, "in.npz"), os.path.join(tmp, "out.npz"))
        env = dict(os.environ, EPIEVO_B200_TIGHT_CAPACITY="1")
        r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=env)
        assert r.returncode == 0, r.stderr[-2000:]
        assert json.loads(r.stdout.strip().splitlines()[-1])["recoveries"] > 0, "the recovery path was not exercised"
        got = np.load(os.path.join(tmp, "out.npz"))
        ref = outs[0][3]
        assert np.array_equal(got["off"], ref.row_off) and np.array_equal(got["jumps"], ref.jumps)
        assert np.array_equal(got["init"], ref.init_state) and np.array_equal(got["J"], outs[0][0])
