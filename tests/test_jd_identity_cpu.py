"""The identity behind suffstats_tile_kernel (epievo_b200/csrc/kernels_stats.cu), restated in plain
Python and checked against the C oracle's row-triple merge (add_sufficient_statistics,
Path.cpp:267-301) -- including histories with MANY exactly equal jump times in neighbouring rows,
which random test data never contains and which pin the tie rule (the higher site jumps first):

  D[trip] = tot * #{triples whose FINAL triplet is trip}
            + sum over events at time t of t * ([triplet before = trip] - [triplet after = trip])
  J[trip] = #{events of the middle row whose triplet before is trip}

with the triplet before an event taken from the initial states and the parity of the neighbours'
jumps ordered before it (left neighbours: strictly earlier; right neighbours: earlier or equal)."""
import numpy as np
import pytest

import cpu_libs
from epievo_b200.paths import PathSet, Tree
from synth import ref_test_tree


def grid_paths(tree, n, seed, levels, p_jump):
    """Jump times on a coarse grid k*T/levels: neighbouring rows tie all the time."""
    rng = np.random.default_rng(seed)
    init = rng.integers(0, 2, (tree.n_nodes, n)).astype(np.uint8)
    init[0] = 0
    lens, chunks = np.zeros((tree.n_nodes, n), np.int64), []
    for v in range(1, tree.n_nodes):
        T = float(tree.branches[v])
        for s in range(n):
            k = np.nonzero(rng.random(levels - 1) < p_jump)[0] + 1
            lens[v, s] = k.size
            chunks.append(k * (T / levels))
    off = np.zeros(tree.n_nodes * n + 1, np.uint64)
    np.cumsum(lens.reshape(-1), out=off[1:].view(np.int64))
    tot = np.array(tree.branches, np.float64); tot[0] = 0.0
    return PathSet(n, tree.n_nodes, init.reshape(-1), off, np.concatenate(chunks), tot, list(tree.node_names))


def jd_by_identity(ps):
    n, J, D = ps.n_sites, np.zeros((ps.n_nodes, 8)), np.zeros((ps.n_nodes, 8))
    lens = ps.row_len()
    init = ps.init_state.reshape(ps.n_nodes, n).astype(np.int64)
    final = init ^ (lens & 1)
    for v in range(1, ps.n_nodes):
        tot = float(ps.tot_time[v])
        C = np.bincount(final[v, :-2] * 4 + final[v, 1:-1] * 2 + final[v, 2:], minlength=8)
        S = np.zeros((8, 3))                       # [triplet before][0: right row, 1: middle, 2: left row jumped]
        rows = [ps.row(v, s) for s in range(n)]

        def state(y, t, ties):                     # state of row y just before our event at t
            r = rows[y]
            return int(init[v, y]) ^ (int(np.count_nonzero(r <= t if ties else r < t)) & 1)

        for s in range(n):
            for k, t in enumerate(rows[s]):
                own = int(init[v, s]) ^ (k & 1)
                if 1 <= s - 1 <= n - 2:            # triple centred on s-1: we are its right row
                    S[state(s - 2, t, False) * 4 + state(s - 1, t, False) * 2 + own, 0] += t
                if 1 <= s <= n - 2:                # centred on s: the middle row
                    tb = state(s - 1, t, False) * 4 + own * 2 + state(s + 1, t, True)
                    S[tb, 1] += t; J[v, tb] += 1
                if 1 <= s + 1 <= n - 2:            # centred on s+1: its left row
                    S[own * 4 + state(s + 1, t, True) * 2 + state(s + 2, t, True), 2] += t
        for trip in range(8):
            D[v, trip] = tot * C[trip] + sum(S[trip, w] - S[trip ^ (1 << w), w] for w in range(3))
    return J, D


@pytest.mark.parametrize("case,n,levels,p", [("testtree", 300, 6, 0.25), ("testtree", 200, 3, 0.6),
                                             ("pair", 400, 4, 0.5), ("bal8", 120, 5, 0.15)])
def test_final_triplet_identity_with_ties(case, n, levels, p):
    tree = {"testtree": ref_test_tree, "pair": lambda: Tree.pair(1.0),
            "bal8": lambda: Tree.balanced(8, 0.05)}[case]()
    ps = grid_paths(tree, n, 17, levels, p)
    o = cpu_libs.oracle()
    o.load_paths(ps)
    Jr, Dr = o.get_suff_stats()
    J, D = jd_by_identity(ps)
    assert np.array_equal(J, Jr)
    assert np.allclose(D, Dr, rtol=1e-12, atol=1e-12)
