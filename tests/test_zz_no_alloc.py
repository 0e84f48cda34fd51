"""epv_suffstats_static (csrc/api_noalloc.cu): the J/D pass with its scratch in a file-scope device array,
for callers that cannot allocate (no counterpart in the reference).
Called through the C-ABI; its output must equal epv_suffstats' on the same resident paths --
J exactly, D to 1e-12 relative (the tolerance of the sharded-vs-single tests; both fold the same integer
sums) -- on two calls in a row: the array keeps the previous call's contents and must be cleared.
Collected last (pytest -x must not hide other tests behind a failure here)."""
import numpy as np
import pytest

from epievo_b200.device import Context, EpvError
from epievo_b200.paths import Tree
from synth import ref_test_tree, synth_paths

pytestmark = pytest.mark.gpu


def test_static_scratch_suffstats_matches_suffstats_twice():
    # three inputs one after the other (a context per tree: the node count of a context is fixed while it
    # holds paths).  The scratch array is ONE per GPU, shared by all of them: stale sums of an earlier
    # call would show up in a later one.  On each input the new entry point is called twice in a row.
    for tree, n, seed, density in ((ref_test_tree(), 5000, 3, 2.0), (Tree.balanced(8, 0.05), 3001, 4, 6.0),
                                   (Tree.balanced(32, 0.05), 700, 5, 3.0)):
        c = Context(0)
        try:
            ps = synth_paths(tree, n, seed, density)
            c.set_tree(tree); c.upload_paths(ps)
            J0, D0 = c.suffstats()
            J1, D1 = c.suffstats_static()
            J2, D2 = c.suffstats_static()                # twice in a row on the same paths
            assert J0[1:].sum() > 0
            for J, D in ((J1, D1), (J2, D2)):
                assert np.array_equal(J, J0)
                assert np.allclose(D, D0, rtol=1e-12, atol=0)
            # time is conserved: sum_t D[b][t] == (n - 2) * branch length
            assert np.allclose(D1[1:].sum(1), (n - 2)*np.asarray(tree.branches)[1:], rtol=1e-9)
            # and on other paths in the same context: upload, static call first this time
            ps2 = synth_paths(tree, n, seed + 100, density*0.5)
            c.upload_paths(ps2)
            J3, D3 = c.suffstats_static()
            J4, D4 = c.suffstats()
            assert np.array_equal(J3, J4) and np.allclose(D3, D4, rtol=1e-12, atol=0)
            assert not np.array_equal(J3, J0)
        finally:
            c.close()


def test_static_scratch_rejects_trees_beyond_its_limit():
    import ctypes as C
    from epievo_b200 import _lib as L
    lib = L.load()
    mx, nb = C.c_uint64(), C.c_uint64()
    assert lib.epv_static_limits(C.byref(mx), C.byref(nb)) == 0
    assert (mx.value, nb.value) == (64, 63*80*8)
    tree = Tree.balanced(64, 0.03)                       # 127 nodes > 64
    c = Context(0)
    try:
        c.set_tree(tree); c.upload_paths(synth_paths(tree, 200, 1, 2.0))
        c.suffstats()                                    # the allocating entry point takes it
        with pytest.raises(EpvError, match="at most 64"):
            c.suffstats_static()
    finally:
        c.close()
