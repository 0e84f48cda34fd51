"""Site sharding with halo exchange: 2/3/4 shards emulated as several contexts on ONE
GPU (records moved by device-to-device copies -- no kernel ever waits on another) must
reproduce the unsharded run bit for bit: paths, cached tri_llh, J/D and accept counts."""
import numpy as np
import pytest

from epievo_b200 import sharded as S
from epievo_b200.device import Context
from epievo_b200.paths import Tree
from epievo_b200.synthetic import simulate_paths
from synth import ref_test_tree

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("case,n,world", [("testtree", 1000, 2), ("testtree", 1001, 3),
                                          ("pair", 500, 4), ("bal8", 302, 2), ("bal32", 100, 3)])
def test_sharded_equals_single(model, case, n, world):
    import torch
    tree = {"testtree": ref_test_tree, "pair": lambda: Tree.pair(1.0),
            "bal8": lambda: Tree.balanced(8, 0.05), "bal32": lambda: Tree.balanced(32, 0.05)}[case]()
    rates, T = model
    ps = simulate_paths(tree, n, rates, T, seed=3)
    dev = torch.device("cuda", 0)
    st = torch.cuda.Stream()             # every context and the record copies share one stream
    torch.cuda.set_stream(st)
    one = Context(0)
    one.set_tree(tree); one.set_model(rates, T); one.upload_paths(ps); one.reset()
    J1, D1, acc1 = one.run_mcmc(2, 3, seed=77, first_sweep=5)
    want = one.download_paths()
    tri1 = one.tri_llh()
    one.close()

    shards = []
    for r in range(world):
        c = Context(0)
        c.set_stream(st.cuda_stream)
        c.set_tree(tree); c.set_model(rates, T)
        sh = S.Shard(c, n, world, r, dev)
        c.upload_paths(S.local_paths(ps, n, world, r)); c.reset()
        shards.append(sh)
    tr = S.LocalExchange(shards)
    J, D, acc = S.run_mcmc(shards, tr, 2, 3, seed=77, first_sweep=5)
    torch.cuda.synchronize()
    parts = [s.ctx.download_paths() for s in shards]
    got = S.gather_paths(parts, n, world)
    assert got.same_as(want)
    assert np.array_equal(J, J1) and np.allclose(D, D1, rtol=1e-12, atol=0) and acc == acc1
    for s in shards:                     # cached tri_llh of owned sites == single-GPU cache
        t = s.ctx.tri_llh()
        assert np.array_equal(t[s.halo_lo:s.halo_lo + (s.hi - s.lo)][1:-1], tri1[s.lo:s.hi][1:-1])
        # halo copies of the neighbours' rows are current
        loc = S.local_paths(want, n, world, s.rank)
        assert parts[s.rank].same_as(loc)
        s.ctx.close()


def test_sharded_with_deferred_overflow_launch(model, monkeypatch):
    """Long passes of a sharded context run three update launches (main with deferral, overflow
    list, halo-aware edge groups); forced on here through EPIEVO_B200_DEFER_MIN, two shards with
    dense rows must still reproduce the unsharded chain without deferral."""
    import torch
    tree = ref_test_tree()
    rates, T = model
    n, world = 900, 2
    from synth import synth_paths
    ps = synth_paths(tree, n, 21, 12.0)
    dev = torch.device("cuda", 0)
    st = torch.cuda.Stream()
    torch.cuda.set_stream(st)
    monkeypatch.setenv("EPIEVO_B200_DEFER_MIN", "1000000000000")
    one = Context(0)
    one.set_tree(tree); one.set_model(rates, T); one.upload_paths(ps); one.reset()
    J1, D1, acc1 = one.run_mcmc(2, 2, seed=13)
    want = one.download_paths()
    one.close()
    monkeypatch.setenv("EPIEVO_B200_DEFER_MIN", "0")
    shards = []
    for r in range(world):
        c = Context(0)
        c.set_stream(st.cuda_stream)
        c.set_tree(tree); c.set_model(rates, T)
        sh = S.Shard(c, n, world, r, dev)
        c.upload_paths(S.local_paths(ps, n, world, r)); c.reset()
        shards.append(sh)
    J, D, acc = S.run_mcmc(shards, S.LocalExchange(shards), 2, 2, seed=13)
    torch.cuda.synchronize()
    got = S.gather_paths([s.ctx.download_paths() for s in shards], n, world)
    for s in shards:
        s.ctx.close()
    assert got.same_as(want)
    assert np.array_equal(J, J1) and np.allclose(D, D1, rtol=1e-12, atol=0) and acc == acc1


@pytest.mark.parametrize("T,start_jumps,tight", [(20.0, 1.0, False), (3.0, 30.0, True)])
def test_long_rows_at_the_shard_ends_through_the_late_capacity_watch(model, monkeypatch, T, start_jumps, tight):
    """The torch driver reads the halo-row capacity check a sweep late (sharded.CapacityWatch).  T = 20 with
    short starting histories: boundary rows grow from ~1 to ~30 jumps in the first sweep -- the capacity set at
    upload must already cover that (it is sized from the branch lengths too).  T = 3 with 30-jump starting rows
    and no headroom at upload (test hook): the rows next to the ends are over half of the record from the start,
    so the watch has to grow it.  Either way the chain equals the unsharded one."""
    import torch
    from synth import synth_paths
    if tight:
        monkeypatch.setenv("EPIEVO_B200_TIGHT_HALO", "1")
    tree = Tree.pair(T)
    rates, Tm = model
    n, world = 600, 3
    ps = synth_paths(tree, n, 9, start_jumps)
    dev = torch.device("cuda", 0)
    st = torch.cuda.Stream()
    torch.cuda.set_stream(st)
    one = Context(0)
    one.set_tree(tree); one.set_model(rates, Tm); one.upload_paths(ps); one.reset()
    J1, D1, acc1 = one.run_mcmc(4, 3, seed=5)
    want = one.download_paths()
    one.close()
    shards = []
    for r in range(world):
        c = Context(0)
        c.set_stream(st.cuda_stream)
        c.set_tree(tree); c.set_model(rates, Tm)
        sh = S.Shard(c, n, world, r, dev)
        c.upload_paths(S.local_paths(ps, n, world, r)); c.reset()
        shards.append(sh)
    cap0 = max(s.ctx.halo_capacity()[0] for s in shards)
    tr = S.LocalExchange(shards)
    assert S.CapacityWatch(shards, tr).ok          # the device-side path, not the synchronising fallback
    J, D, acc = S.run_mcmc(shards, tr, 4, 3, seed=5)
    torch.cuda.synchronize()
    got = S.gather_paths([s.ctx.download_paths() for s in shards], n, world)
    assert got.same_as(want)
    assert np.array_equal(J, J1) and np.array_equal(D, D1) and acc == acc1
    caps = {s.ctx.halo_capacity()[0] for s in shards}
    assert len(caps) == 1                          # the same record size everywhere
    if tight:
        assert caps.pop() > cap0                   # grown by the watch
    else:
        assert cap0 >= 6 * T                       # sized from the branch length at upload
    for s in shards:
        s.ctx.close()
