"""The compact exchange format (u16 row codes + jump times, include/epievo_b200.h
epv_upload_paths_compact / epv_download_paths_compact) is the CSR format with the per-row
arrays packed: both must place identical paths on the device and read identical paths
back -- whole problems, shards with halos, before and after sweeps, and when the context is
re-used for another upload of the same or another shape."""
import numpy as np
import pytest

from epievo_b200 import sharded as S
from epievo_b200.device import Context, EpvError
from epievo_b200.paths import PathSet, Tree
from epievo_b200.synthetic import simulate_paths
from synth import ref_test_tree

pytestmark = pytest.mark.gpu

TREES = {"testtree": ref_test_tree, "pair": lambda: Tree.pair(1.0),
         "bal8": lambda: Tree.balanced(8, 0.05), "bal32": lambda: Tree.balanced(32, 0.05),
         "bal64": lambda: Tree.balanced(64, 0.02)}


def test_row_code_round_trip_host(model):
    rates, T = model
    ps = simulate_paths(ref_test_tree(), 257, rates, T, seed=1)
    back = PathSet.from_row_code(ps.n_sites, ps.n_nodes, ps.row_code(), ps.jumps, ps.tot_time)
    assert back.same_as(ps)


@pytest.mark.parametrize("case,n", [("testtree", 3), ("testtree", 1023), ("testtree", 1024),
                                    ("testtree", 4100), ("pair", 2500), ("bal8", 3000),
                                    ("bal32", 700), ("bal64", 130)])
def test_compact_equals_csr(model, case, n):
    rates, T = model
    tree = TREES[case]()
    ps = simulate_paths(tree, n, rates, T, seed=11)
    a, b = Context(0), Context(0)
    try:
        for c in (a, b):
            c.set_tree(tree); c.set_model(rates, T)
        a.upload_paths(ps)
        b.upload_paths_compact(ps)
        assert b.download_paths().same_as(ps)            # compact in, CSR out
        assert a.download_paths_compact().same_as(ps)    # CSR in, compact out
        a.reset(); b.reset()
        assert np.array_equal(a.tri_llh(), b.tri_llh())
        Ja, Da = a.suffstats(); Jb, Db = b.suffstats()
        assert np.array_equal(Ja, Jb) and np.array_equal(Da, Db)
        if n >= 6:
            ra = a.run_mcmc(2, 2, seed=5)
            rb = b.run_mcmc(2, 2, seed=5)
            assert np.array_equal(ra[0], rb[0]) and np.array_equal(ra[1], rb[1]) and ra[2] == rb[2]
            out = a.download_paths()
            assert b.download_paths_compact().same_as(out)
            # re-use: same shape again (buffers kept), then another shape
            b.upload_paths_compact(out)
            assert b.download_paths().same_as(out)
            small = simulate_paths(tree, max(3, n // 3), rates, T, seed=12)
            b.upload_paths_compact(small)
            assert b.download_paths_compact().same_as(small)
            a.upload_paths(small)
            assert a.download_paths().same_as(small)
    finally:
        a.close(); b.close()


@pytest.mark.parametrize("world", [2, 3])
def test_compact_shards_with_halos(model, world):
    import torch
    rates, T = model
    tree = Tree.balanced(8, 0.05)
    n = 1500
    ps = simulate_paths(tree, n, rates, T, seed=4)
    dev = torch.device("cuda", 0)
    for r in range(world):
        loc = S.local_paths(ps, n, world, r)
        a, b = Context(0), Context(0)
        try:
            for c in (a, b):
                c.set_tree(tree); c.set_model(rates, T)
                S.Shard(c, n, world, r, dev)
            a.upload_paths(loc); b.upload_paths_compact(loc)
            assert b.download_paths().same_as(loc)
            assert b.download_paths_compact().same_as(loc)
            a.reset(); b.reset()
            assert np.array_equal(a.tri_llh(), b.tri_llh())
            for col in range(3):
                a.colour_pass(col, 9, 0); b.colour_pass(col, 9, 0)
            assert a.download_paths().same_as(b.download_paths_compact())
        finally:
            a.close(); b.close()


def test_compact_rejects_bad_input(model):
    rates, T = model
    tree = ref_test_tree()
    ps = simulate_paths(tree, 100, rates, T, seed=2)
    c = Context(0)
    try:
        c.set_tree(tree); c.set_model(rates, T)
        code = ps.row_code()
        j = ps.jumps
        with pytest.raises(EpvError, match="add up"):
            c.upload_compact_raw(ps.n_sites, ps.n_nodes, code.ctypes.data, j.ctypes.data, len(j) + 1,
                                 ps.tot_time)
        c.upload_paths_compact(ps)
        small = np.zeros(1)
        with pytest.raises(EpvError, match="too small"):
            c.download_compact_raw(code.ctypes.data, small.ctypes.data, 1)
    finally:
        c.close()
