"""Exact posterior draws of whole-tree histories on a tiny problem, by forward simulation of the
context-dependent process along every branch + rejection on the leaf sequences -- the design of
the reference harness (end_cond_forward_simulation, /root/reference/src/harnesses/MCMC_test.cpp:192-216)
extended from one branch to a tree.  TEST INFRASTRUCTURE (numpy restatement of the forward process
the reference simulates in prog/epievo_sim.cpp:102-152 / TripletSampler.cpp: global holding time from
the sum of the per-site triplet rates, then a site proportional to its rate; the two end sites never
change, TripletSampler.cpp:58-60).  Pinned to the reference harness's own exact draws on the
single-branch case by tests/test_exact_posterior_cpu.py.

Vectorised over replicates: R independent copies of the whole sequence advance one event at a time.
"""
from __future__ import annotations

from typing import List, Tuple

import numpy as np

from epievo_b200.paths import PathSet, Tree


def _triplets(x: np.ndarray) -> np.ndarray:
    """[R, n] states -> [R, n-2] triplet index l*4+m*2+r of every interior site."""
    return (x[:, :-2].astype(np.int64) << 2) | (x[:, 1:-1] << 1) | x[:, 2:]


def _counts(trip: np.ndarray) -> np.ndarray:
    return (trip[:, :, None] == np.arange(8)[None, None, :]).sum(1).astype(np.float64)


def simulate_branch(x0: np.ndarray, T: float, rates: np.ndarray, rng: np.random.Generator):
    """All replicates forward over one branch.  Returns end states [R, n], J [R, 8], D [R, 8]
    (the triple statistics of add_sufficient_statistics summed over the interior sites)."""
    x = x0.copy()
    R = x.shape[0]
    J, D = np.zeros((R, 8)), np.zeros((R, 8))
    t = np.zeros(R)
    live = np.arange(R)
    while live.size:
        xs = x[live]
        trip = _triplets(xs)
        lam = rates[trip]                                   # [r, n-2]
        tot = lam.sum(1)
        dt = rng.exponential(1.0, live.size)/tot
        cnt = _counts(trip)
        over = t[live] + dt >= T
        D[live] += cnt*np.where(over, T - t[live], dt)[:, None]
        t[live] += dt
        keep = ~over
        if keep.any():
            idx = live[keep]
            cum = np.cumsum(lam[keep], 1)
            u = rng.random(idx.size)*cum[:, -1]
            site = (cum <= u[:, None]).sum(1)               # interior index 0..n-3
            site = np.minimum(site, lam.shape[1] - 1)
            J[idx, trip[keep, site]] += 1.0
            x[idx, site + 1] ^= 1
        live = live[keep]
    return x, J, D


def exact_draws(tree: Tree, root: np.ndarray, leaves: dict, rates, n_draws: int, seed: int,
                batch: int = 200000, max_batches: int = 4000) -> Tuple[np.ndarray, float]:
    """[n_draws, n_nodes, 16] per-branch (J[8], D[8]) of exact posterior draws given the root
    sequence and the leaf sequences (leaves: node -> sequence); acceptance rate."""
    rng = np.random.default_rng(seed)
    rates = np.asarray(rates, np.float64)
    n_nodes = tree.n_nodes
    out: List[np.ndarray] = []
    got = tried = 0
    for _ in range(max_batches):
        end = {0: np.repeat(root[None, :].astype(np.uint8), batch, 0)}
        ok = np.ones(batch, bool)
        stats = np.zeros((batch, n_nodes, 16))
        for v in range(1, n_nodes):
            e, J, D = simulate_branch(end[int(tree.parent_ids[v])], float(tree.branches[v]), rates, rng)
            end[v] = e
            stats[:, v, :8] = J; stats[:, v, 8:] = D
            if v in leaves:
                ok &= (e == np.asarray(leaves[v], np.uint8)[None, :]).all(1)
        out.append(stats[ok]); got += int(ok.sum()); tried += batch
        if got >= n_draws:
            break
    assert got >= n_draws, "acceptance too low: %d of %d" % (got, tried)
    return np.concatenate(out)[:n_draws], got/tried


def make_problem(tree: Tree, n_sites: int, rates, T, seed: int):
    """Root sequence from the horizontal chain, leaf data from one forward simulation."""
    rng = np.random.default_rng(seed)
    root = np.zeros(n_sites, np.uint8)
    T00, T11 = float(T[0]), float(T[3])
    root[0] = rng.random() < (1 - T00)/(2 - T00 - T11)
    for s in range(1, n_sites):
        stay = T11 if root[s - 1] else T00
        root[s] = root[s - 1] if rng.random() < stay else 1 - root[s - 1]
    end = {0: root[None, :]}
    for v in range(1, tree.n_nodes):
        end[v], _, _ = simulate_branch(end[int(tree.parent_ids[v])], float(tree.branches[v]),
                                       np.asarray(rates, np.float64), rng)
    leaves = {v: end[v][0] for v in range(1, tree.n_nodes) if tree.is_leaf(v)}
    return root, leaves


def initial_paths(tree: Tree, root: np.ndarray, leaves: dict, seed: int) -> PathSet:
    """A valid starting history: internal nodes carry the root sequence, every branch one uniformly
    placed jump where its two ends differ (initialize_paths, MCMC_test.cpp:219-237)."""
    rng = np.random.default_rng(seed)
    n = len(root)
    seq = {0: root}
    for v in range(1, tree.n_nodes):
        seq[v] = np.asarray(leaves[v], np.uint8) if v in leaves else root
    rows = [[[] for _ in range(n)]]
    init = [np.zeros(n, np.uint8)]
    for v in range(1, tree.n_nodes):
        a = seq[int(tree.parent_ids[v])]
        Tv = float(tree.branches[v])
        rows.append([[rng.random()*Tv] if a[s] != seq[v][s] else [] for s in range(n)])
        init.append(a)
    tot = np.array(tree.branches, np.float64); tot[0] = 0.0
    return PathSet.from_rows(n, tree.n_nodes, init, rows, tot, list(tree.node_names))
