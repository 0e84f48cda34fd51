"""TEST INFRASTRUCTURE ONLY -- numpy restatement of the reference's site-independent E-step
(/root/reference/src/libepievo/IndepSite.cpp), used by tests/ as the checker of
epv_indep_expectation / epv_indep_suffstats / epv_indep_update_paths.  Nothing under
epievo_b200/ imports this file.  Pinned against the unmodified reference (oracle/_ref) by
tests/test_indep_cpu.py.  Vectorised over sites; per-node loops follow the reference."""
import numpy as np


def trans_prob_mat(r0, r1, T):
    """continuous_time_trans_prob_mat, ContinuousTimeMarkovModel.cpp:142-161 (h = 1/exp(+x))."""
    h = 1.0 / np.exp(T * (r0 + r1))
    s = r0 + r1
    P = np.zeros((2, 2))
    P[0, 0] = (r0 * h + r1) / s
    P[0, 1] = 1.0 - P[0, 0]
    P[1, 1] = (r0 + r1 * h) / s
    P[1, 0] = 1.0 - P[1, 1]
    return P


def expectation_J(r0, r1, T):
    """ContinuousTimeMarkovModel.cpp:167-195."""
    s, p, d, e = r0 + r1, r0 * r1, r1 - r0, np.exp(-(r0 + r1) * T)
    J0, J1 = np.zeros((2, 2)), np.zeros((2, 2))
    C1 = d * (1 - e) / s
    J0[0, 0] = p * (T * (r1 - r0 * e) - C1) / (s * (r1 + r0 * e)); J1[0, 0] = J0[0, 0]
    J0[1, 1] = p * (T * (r0 - r1 * e) + C1) / (s * (r0 + r1 * e)); J1[1, 1] = J0[1, 1]
    C2 = p * T * (1 + e) / (s * (1 - e))
    C3 = (r0 * r0 + r1 * r1) / (s * s)
    C4 = (2 * p) / (s * s)
    J0[0, 1] = C2 + C3; J1[0, 1] = C2 - C4
    J0[1, 0] = J1[0, 1]; J1[1, 0] = J0[0, 1]
    return J0, J1


def expectation_D(r0, r1, T):
    """ContinuousTimeMarkovModel.cpp:198-226."""
    r00, r11, s, p = r0 * r0, r1 * r1, r0 + r1, r0 * r1
    e = np.exp(-s * T)
    D0, D1 = np.zeros((2, 2)), np.zeros((2, 2))
    C1 = 2 * p * (1 - e) / s
    D0[0, 0] = ((r11 + r00 * e) * T + C1) / (s * (r1 + r0 * e)); D1[0, 0] = T - D0[0, 0]
    D1[1, 1] = ((r00 + r11 * e) * T + C1) / (s * (r0 + r1 * e)); D0[1, 1] = T - D1[1, 1]
    C2 = (p - r00) * (1 - e) / s
    D1[0, 1] = ((r00 - p * e) * T + C2) / (s * (r0 - r0 * e)); D0[0, 1] = T - D1[0, 1]
    C3 = (p - r11) * (1 - e) / s
    D0[1, 0] = ((r11 - p * e) * T + C3) / (s * (r1 - r1 * e)); D1[1, 0] = T - D0[1, 0]
    return D0, D1


def upward(tree, ps, rates):
    """upward_process, IndepSite.cpp:53-89, for all sites at once: q[node] is [n_sites, 2]."""
    n = tree.n_nodes
    end = ps.end_state().astype(np.int64)                  # [n_nodes, n_sites]
    q = [np.ones((ps.n_sites, 2)) for _ in range(n)]
    p = [None] * n
    for v in range(n - 1, -1, -1):
        if tree.is_leaf(v):
            q[v] = np.stack([(end[v] == 0).astype(float), (end[v] == 1).astype(float)], axis=1)
        if v > 0:
            P = trans_prob_mat(rates[0], rates[1], float(tree.branches[v]))
            p[v] = q[v] @ P.T                                # p = P q
            par = int(tree.parent_ids[v])
            q[par] = q[par] * p[v]
    return q, p


def expectation_sufficient_statistics(tree, ps, rates, root=None):
    """IndepSite.cpp:221-239 (+ downward_process :157-171, weighted_J_D_branch :123-154,
    joint_post :103-120, root_post_prob0 :92-99).  Returns J, D [n_nodes, 2], sums over sites.
    root: optional per-site root states to condition on instead of marginalising the root
    (what sampling_downward, :176-214, does: it keeps the root state) -- test use only."""
    n = tree.n_nodes
    q, p = upward(tree, ps, rates)
    pi0 = rates[1] / (rates[0] + rates[1])
    a, b = pi0 * q[0][:, 0], (1 - pi0) * q[0][:, 1]
    p0m = [None] * n
    p0m[0] = a / (a + b) if root is None else (np.asarray(root) == 0).astype(float)
    J, D = np.zeros((n, 2)), np.zeros((n, 2))
    for v in range(1, n):
        T = float(tree.branches[v])
        P = trans_prob_mat(rates[0], rates[1], T)
        pu = p0m[int(tree.parent_ids[v])]
        pj = np.zeros((ps.n_sites, 2, 2))
        pj[:, 0, 0] = P[0, 0] * q[v][:, 0] * pu / p[v][:, 0]
        pj[:, 0, 1] = P[0, 1] * q[v][:, 1] * pu / p[v][:, 0]
        pj[:, 1, 0] = P[1, 0] * q[v][:, 0] * (1 - pu) / p[v][:, 1]
        pj[:, 1, 1] = P[1, 1] * q[v][:, 1] * (1 - pu) / p[v][:, 1]
        pj /= pj.sum(axis=(1, 2), keepdims=True)
        p0m[v] = pj[:, 0, 0] + pj[:, 1, 0]
        J0, J1 = expectation_J(rates[0], rates[1], T)
        D0, D1 = expectation_D(rates[0], rates[1], T)
        J[v, 0] = (pj * J0).sum(); J[v, 1] = (pj * J1).sum()
        D[v, 0] = (pj * D0).sum(); D[v, 1] = (pj * D1).sum()
    return J, D


def compute_sufficient_statistics(ps):
    """IndepSite.cpp:266-303, as SUMS over sites (the reference divides by n_sites)."""
    n, ns = ps.n_nodes, ps.n_sites
    J, D = np.zeros((n, 2)), np.zeros((n, 2))
    off = ps.row_off.astype(np.int64)
    for b in range(1, n):
        T = float(ps.tot_time[b])
        for s in range(ns):
            r = b * ns + s
            prev, time = int(ps.init_state[r]), 0.0
            for t in ps.jumps[off[r]:off[r + 1]]:
                J[b, prev] += 1
                D[b, prev] += t - time
                prev, time = 1 - prev, t
            D[b, prev] += T - time
    return J, D
