"""TEST INFRASTRUCTURE ONLY -- CPU restatement (numpy) of the forward-simulation algorithm of
epievo_b200/csrc/api_sim.cu; nothing under epievo_b200/ imports it.

What it restates: the reference's epievo_sim draws one event of the whole sequence at a time
(prog/epievo_sim.cpp:102-152 sample_jump: holding time from sum c_ijk*lambda_ijk, triplet class,
position inside the class, TripletSampler.cpp).  The device algorithm reaches the same law by
  * thinning: candidate events of rate Lambda = max lambda per interior site, a candidate at
    (s, t) flips the site with probability lambda[x_{s-1}(t) x_s(t) x_{s+1}(t)] / Lambda; the two
    end sites never change (TripletSampler.cpp:58-60);
  * order: a candidate is decided once neither neighbour has an undecided candidate before it,
    i.e. when its time is a local minimum of the per-site next-candidate times (ties by site);
    all local minima are decided together, round after round.
The random stream is numpy's, not the device's Philox, so this pins the ALGORITHM's law, through
the same statistics and z-test as the device test (tests/golden/sim_stats.json, made from the
reference binaries by tests/golden/make_sim_golden.py).  Parity status: pinned statistically.
"""
import numpy as np


def root_sequence(n, T, rng):
    """sample_state_sequence (EpiEvoModel.cpp:282-298): stationary start, then stay with
    probability T[x][x].  T is row-major 2x2."""
    T00, T11 = T[0], T[3]
    pi1 = (1.0 - T00) / (2.0 - T11 - T00)
    u = rng.random(n)
    x = np.empty(n, np.uint8)
    x[0] = u[0] < pi1
    # the chain flips whenever u > stay-probability of the current state: sequential by nature
    for i in range(1, n):
        p = T11 if x[i - 1] else T00
        x[i] = x[i - 1] if u[i] <= p else 1 - x[i - 1]
    return x


def simulate_branch(x0, T_branch, rates, rng):
    """One branch: returns (end states, jump sites, jump times), jumps grouped by site in time
    order (each site's candidates are decided in increasing time)."""
    rates = np.asarray(rates, np.float64)
    lam = float(rates.max())
    n = x0.size
    x = x0.astype(np.int64).copy()
    tn = rng.exponential(1.0 / lam, n)
    tn[0] = tn[-1] = np.inf
    tn[tn >= T_branch] = np.inf
    sites, times = [], []
    while np.isfinite(tn).any():
        left = np.concatenate(([np.inf], tn[:-1]))
        right = np.concatenate((tn[1:], [np.inf]))
        idx = np.nonzero(np.isfinite(tn) & (tn < left) & (tn <= right))[0]
        ctx = x[idx - 1] * 4 + x[idx] * 2 + x[idx + 1]
        acc = rng.random(idx.size) * lam < rates[ctx]
        sites.append(idx[acc]); times.append(tn[idx][acc])
        x[idx[acc]] ^= 1
        nxt = tn[idx] + rng.exponential(1.0 / lam, idx.size)
        nxt[nxt >= T_branch] = np.inf
        tn[idx] = nxt
    sites = np.concatenate(sites) if sites else np.zeros(0, np.int64)
    times = np.concatenate(times) if times else np.zeros(0)
    order = np.lexsort((times, sites))
    return x.astype(np.uint8), sites[order], times[order]


def simulate_paths(tree, n_sites, rates, T, seed, root=None):
    """Histories down the tree in the C-ABI exchange format, as a PathSet."""
    from epievo_b200.paths import PathSet
    rng = np.random.default_rng(seed)
    n_nodes = tree.n_nodes
    init = np.zeros((n_nodes, n_sites), np.uint8)
    end = np.zeros((n_nodes, n_sites), np.uint8)
    end[0] = root_sequence(n_sites, T, rng) if root is None else np.asarray(root, np.uint8)
    lens = np.zeros((n_nodes, n_sites), np.int64)
    chunks = []
    for v in range(1, n_nodes):
        p = int(tree.parent_ids[v])
        init[v] = end[p]
        end[v], js, jt = simulate_branch(end[p], float(tree.branches[v]), rates, rng)
        lens[v] = np.bincount(js, minlength=n_sites)
        chunks.append(jt)
    off = np.zeros(n_nodes * n_sites + 1, np.uint64)
    np.cumsum(lens.reshape(-1), out=off[1:].view(np.int64))
    tot = np.array(tree.branches, np.float64)
    tot[0] = 0.0
    return PathSet(n_sites, n_nodes, init.reshape(-1), off, np.concatenate(chunks) if chunks else np.zeros(0),
                   tot, list(tree.node_names))
