"""Synthetic inputs that need neither the reference nor a GPU: valid (tree-consistent)
histories with a chosen jump density, the test tree / parameters of the reference's
test/ directory, and helpers shared by the test modules."""
from __future__ import annotations

import numpy as np

from epievo_b200.paths import PathSet, Tree, parse_newick

TEST_TREE_NWK = "((C:0.03,D:0.06)E:0.02,F:0.1)G:0.0;"        # reference test/tree.nwk
TEST_PARAMS = (0.844912, 0.893359, -0.8, -1.8)                # reference test/test.param


def ref_test_tree() -> Tree:
    return parse_newick(TEST_TREE_NWK)


def synth_paths(tree: Tree, n_sites: int, seed: int, density: float = 1.0,
                stay: float = 0.87) -> PathSet:
    """Random histories: root sequence from a 2-state Markov chain along the sites,
    Poisson(density * branch length) uniformly placed jumps per row, every branch
    starting in its parent's end state."""
    rng = np.random.default_rng(seed)
    n_nodes = tree.n_nodes
    root = np.zeros(n_sites, np.uint8)
    root[0] = rng.random() < 0.5
    flips = rng.random(n_sites) > stay
    for s in range(1, n_sites):
        root[s] = root[s - 1] ^ flips[s]
    init = np.zeros((n_nodes, n_sites), np.uint8)
    end = np.zeros((n_nodes, n_sites), np.uint8)
    end[0] = root
    lens = np.zeros((n_nodes, n_sites), np.int64)
    chunks = []
    for b in range(1, n_nodes):
        T = float(tree.branches[b])
        init[b] = end[int(tree.parent_ids[b])]
        lens[b] = rng.poisson(density * T, n_sites)
        end[b] = (init[b] + lens[b]) & 1
        tot = int(lens[b].sum())
        t = rng.random(tot) * T
        # sort within rows: sort by (row id, time)
        rid = np.repeat(np.arange(n_sites), lens[b])
        order = np.lexsort((t, rid))
        chunks.append(t[order])
    off = np.zeros(n_nodes * n_sites + 1, np.uint64)
    np.cumsum(lens.reshape(-1), out=off[1:].view(np.int64))
    jumps = np.concatenate(chunks) if chunks else np.zeros(0)
    tot_time = np.array(tree.branches, np.float64)
    tot_time[0] = 0.0
    return PathSet(n_sites, n_nodes, init.reshape(-1), off, jumps, tot_time,
                   list(tree.node_names))


def rel_close(a, b, rtol=1e-9, atol=0.0) -> bool:
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    return bool(np.all(np.abs(a - b) <= atol + rtol * np.maximum(np.abs(a), np.abs(b))))
