"""Race freedom of a stride-3 colour class, checked dynamically (compute-sanitizer's racecheck is closed on
this GPU pool): (1) write sets -- after ONE colour pass only sites of that colour have changed, the two end
sites never; (2) read sets -- the same passes under another schedule (EPIEVO_B200_REVERSE: thread i of the
update launch takes group g_count-1-i, so every site is updated by another warp at another time) give the
same chain bit for bit: an update that read anything a concurrent update writes would see other values.
Reference semantics: SingleSiteSampler.cpp:482-536 reads sites s-2..s+2 and writes site s and
tri_llh[s-1..s+1]; same-colour sites are 3 apart."""
import numpy as np
import pytest

from epievo_b200.device import Context
from epievo_b200.paths import Tree
from synth import ref_test_tree, synth_paths

pytestmark = pytest.mark.gpu


def rows_by_site(ps):
    """[n_sites] tuple per site: (init bits of all nodes, all jump times) for comparison."""
    out = []
    for s in range(ps.n_sites):
        out.append((tuple(ps.init(v, s) for v in range(1, ps.n_nodes)),
                    tuple(np.concatenate([ps.row(v, s) for v in range(1, ps.n_nodes)]).tolist())))
    return out


@pytest.mark.parametrize("case", ["testtree", "bal8"])
def test_a_colour_pass_writes_only_its_colour(ctx, model, case):
    tree = ref_test_tree() if case == "testtree" else Tree.balanced(8, 0.05)
    n = 900
    ps = synth_paths(tree, n, 3, 4.0)
    ctx.set_tree(tree); ctx.set_model(*model); ctx.upload_paths(ps); ctx.reset()
    before = rows_by_site(ctx.download_paths())
    for k, colour in enumerate((1, 0, 2, 2, 1)):
        ctx.colour_pass(colour, 17, k)
        after = rows_by_site(ctx.download_paths())
        changed = [s for s in range(n) if after[s] != before[s]]
        assert changed, "no update accepted?"
        assert all(s % 3 == colour for s in changed), [s for s in changed if s % 3 != colour][:5]
        assert 0 not in changed and n - 1 not in changed
        assert len(changed) > 0.3*(n/3)
        before = after


@pytest.mark.parametrize("case,n,density", [("testtree", 40000, 1.0), ("bal8", 20000, 3.0), ("pair", 30000, 1.0),
                                            ("bal32", 6000, 2.0)])
def test_reversed_schedule_gives_the_same_chain(model, case, n, density, monkeypatch):
    tree = {"testtree": ref_test_tree, "pair": lambda: Tree.pair(1.0), "bal8": lambda: Tree.balanced(8, 0.05),
            "bal32": lambda: Tree.balanced(32, 0.05)}[case]()
    ps = synth_paths(tree, n, 8, density)
    out = []
    for rev in (False, True):
        if rev:
            monkeypatch.setenv("EPIEVO_B200_REVERSE", "1")
        else:
            monkeypatch.delenv("EPIEVO_B200_REVERSE", raising=False)
        c = Context(0)
        try:
            c.set_tree(tree); c.set_model(*model); c.upload_paths(ps); c.reset()
            J, D, acc = c.run_mcmc(3, 2, seed=23)
            out.append((J, D, acc, c.download_paths(), c.tri_llh()))
        finally:
            c.close()
    a, b = out
    assert a[2] == b[2] and np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    assert a[3].same_as(b[3]) and np.array_equal(a[4], b[4])
