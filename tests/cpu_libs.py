"""ctypes bindings to the two CPU checkers (TEST INFRASTRUCTURE ONLY):

* ``oracle/liboracle.so``          -- plain-C restatement (prefix ``orc_``)
* ``oracle/_ref/libepievo_ref.so`` -- the unmodified reference sources + driver
                                      (prefix ``ref_``); present only where
                                      oracle/Makefile could see /root/reference
                                      (or where the prebuilt file travelled).

Both export the same entry points, so one wrapper class serves both.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from typing import Optional, Tuple

import numpy as np

from epievo_b200.paths import PathSet, Tree

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
REF_BIN = os.path.join(ORACLE_DIR, "_ref", "bin")

_dp = C.POINTER(C.c_double)
_u64p = C.POINTER(C.c_uint64)
_u8p = C.POINTER(C.c_uint8)


def _d(a: np.ndarray):
    assert a.dtype == np.float64 and a.flags.c_contiguous
    return a.ctypes.data_as(_dp)


def _u64(a: np.ndarray):
    assert a.dtype == np.uint64 and a.flags.c_contiguous
    return a.ctypes.data_as(_u64p)


def _u8(a: np.ndarray):
    assert a.dtype == np.uint8 and a.flags.c_contiguous
    return a.ctypes.data_as(_u8p)


class CpuLib:
    def __init__(self, path: str, prefix: str):
        self.lib = C.CDLL(path)
        self.prefix = prefix
        self.kind = "reference" if prefix == "ref_" else "port"
        f = self._f
        f("trans_prob_mat", None, [C.c_double] * 3 + [_dp])
        f("trans_prob", C.c_double, [C.c_double] * 3 + [C.c_int] * 2)
        f("expectations", None, [C.c_double] * 3 + [_dp] * 4)
        f("segment_info", C.c_uint64,
          [_dp, C.c_int, C.c_uint64, _dp, C.c_int, C.c_uint64, _dp, C.c_double,
           _dp, _dp, _u64p, _u64p, _dp])
        f("add_suff_stats", None,
          [C.c_int, C.c_uint64, _dp] * 3 + [C.c_double, _dp, _dp])
        f("load_paths", None, [C.c_uint64, C.c_uint64, _u8p, _u64p, _dp, _dp])
        f("total_jumps", C.c_uint64, [])
        f("store_paths", None, [_u8p, _u64p, _dp, _dp])
        f("get_suff_stats", None, [_dp, _dp])
        f("tri_llh", None, [_dp, _dp, _dp])
        f("site_debug", C.c_uint64,
          [C.c_uint64, _u64p, _u64p, _dp, _dp, _dp, C.c_uint64, C.c_uint64,
           _u64p, _dp, _dp, _u64p, _u64p, _dp, _dp, _dp, _dp, _dp])
        f("seed", None, [C.c_uint64])
        f("run_mcmc", C.c_double,
          [C.c_uint64, _u64p, _u64p, _dp, _dp, _dp, C.c_uint64, C.c_uint64, C.c_int,
           _dp, _dp, _dp, _dp])
        f("time_suff_stats", C.c_double, [C.c_uint64, _dp, _dp])
        f("end_cond_sample", C.c_uint64,
          [C.c_double, C.c_double, C.c_int, C.c_int, C.c_double, C.c_uint64, _dp])
        f("model_from_params", None, [C.c_double] * 4 + [C.c_int] + [_dp] * 4)
        f("model_rebuild", None, [_dp] * 4)
        f("rate_scaling_factor", C.c_double, [_dp])
        f("estimate_rates", C.c_double, [C.c_uint64, _dp, _dp, C.c_double, _dp, _dp, _dp])
        f("estimate_rates_and_branches", C.c_double,
          [C.c_uint64, _dp, _dp, C.c_double, _dp, _dp, _dp, _dp])
        f("set_one_change", None, [_dp, C.c_uint64, _dp])
        f("scale_jump_times", None, [C.c_uint64, _dp])
        if prefix == "ref_":
            f("indep_load", None, [C.c_uint64, C.c_uint64, _u8p, _u64p, _dp, _dp])
            f("indep_total_jumps", C.c_uint64, [])
            f("indep_store", None, [_u8p, _u64p, _dp, _dp])
            f("indep_suffstats", None, [C.c_uint64, _dp, _dp])
            f("indep_expectation", None, [C.c_uint64, _u64p, _u64p, _dp, _dp, _dp, _dp])
            f("indep_update_paths", None, [C.c_uint64, _u64p, _u64p, _dp, _dp, C.c_uint64])
            f("indep_estimate", None, [C.c_uint64, _u64p, _u64p, _dp, _dp, _dp, _dp, C.c_int])
            f("read_paths_file", C.c_double, [C.c_char_p, _u64p, _u64p])
            f("write_paths_file", C.c_double, [C.c_char_p, C.c_uint64, C.POINTER(C.c_char_p)])
        if prefix == "orc_":
            f("log_likelihood", C.c_double, [_dp, _dp, _dp])
            f("gradient", None, [_dp] * 4)
        self._n_sites = 0
        self._n_nodes = 0

    def _f(self, name, restype, argtypes):
        fn = getattr(self.lib, self.prefix + name)
        fn.restype = restype
        fn.argtypes = argtypes
        setattr(self, "_" + name, fn)

    # -- scalar kernels -----------------------------------------------------------
    def trans_prob_mat(self, r0, r1, t) -> np.ndarray:
        P = np.zeros(4)
        self._trans_prob_mat(r0, r1, t, _d(P))
        return P

    def trans_prob(self, r0, r1, t, a, b) -> float:
        return self._trans_prob(r0, r1, t, a, b)

    def expectations(self, r0, r1, T):
        out = [np.zeros(4) for _ in range(4)]
        self._expectations(r0, r1, T, *[_d(x) for x in out])
        return out  # J0, J1, D0, D1 (row-major 2x2 over (start, end))

    def segment_info(self, rates, l_init, lj, r_init, rj, tot_time):
        lj = np.ascontiguousarray(lj, np.float64)
        rj = np.ascontiguousarray(rj, np.float64)
        cap = len(lj) + len(rj) + 1
        r0, r1, ln = np.zeros(cap), np.zeros(cap), np.zeros(cap)
        t0, t1 = np.zeros(cap, np.uint64), np.zeros(cap, np.uint64)
        n = self._segment_info(_d(np.ascontiguousarray(rates)), l_init, len(lj), _d(lj),
                               r_init, len(rj), _d(rj), tot_time,
                               _d(r0), _d(r1), _u64(t0), _u64(t1), _d(ln))
        assert n == cap
        return r0, r1, t0, t1, ln

    def add_suff_stats(self, l, m, r, tot_time):
        """l, m, r are (init_state, jumps) pairs; returns (J[8], D[8])."""
        J, D = np.zeros(8), np.zeros(8)
        args = []
        keep = []
        for init, j in (l, m, r):
            j = np.ascontiguousarray(j, np.float64)
            keep.append(j)
            args += [int(init), len(j), _d(j)]
        self._add_suff_stats(*args, tot_time, _d(J), _d(D))
        return J, D

    # -- path container -----------------------------------------------------------
    def load_paths(self, ps: PathSet) -> None:
        self._n_sites, self._n_nodes = ps.n_sites, ps.n_nodes
        self._load_paths(ps.n_sites, ps.n_nodes, _u8(ps.init_state), _u64(ps.row_off),
                         _d(ps.jumps if len(ps.jumps) else np.zeros(1)), _d(ps.tot_time))

    def store_paths(self) -> PathSet:
        n_rows = self._n_sites * self._n_nodes
        init = np.zeros(n_rows, np.uint8)
        off = np.zeros(n_rows + 1, np.uint64)
        jumps = np.zeros(max(1, int(self._total_jumps())))
        tot = np.zeros(self._n_nodes)
        self._store_paths(_u8(init), _u64(off), _d(jumps), _d(tot))
        return PathSet(self._n_sites, self._n_nodes, init, off,
                       jumps[:int(off[-1])].copy(), tot)

    # -- IndepSite.cpp through the reference build -------------------------------------------
    def indep_load(self, ps: PathSet) -> None:
        self._i_sites, self._i_nodes = ps.n_sites, ps.n_nodes
        self._indep_load(ps.n_sites, ps.n_nodes, _u8(ps.init_state), _u64(ps.row_off),
                         _d(ps.jumps if len(ps.jumps) else np.zeros(1)), _d(ps.tot_time))

    def indep_store(self) -> PathSet:
        n_rows = self._i_sites * self._i_nodes
        init, off = np.zeros(n_rows, np.uint8), np.zeros(n_rows + 1, np.uint64)
        jumps, tot = np.zeros(max(1, int(self._indep_total_jumps()))), np.zeros(self._i_nodes)
        self._indep_store(_u8(init), _u64(off), _d(jumps), _d(tot))
        return PathSet(self._i_sites, self._i_nodes, init, off, jumps[:int(off[-1])].copy(), tot)

    def indep_suffstats(self):
        """compute_sufficient_statistics: AVERAGES over sites, [n_nodes, 2]."""
        J, D = np.zeros(self._i_nodes * 2), np.zeros(self._i_nodes * 2)
        self._indep_suffstats(self._i_nodes, _d(J), _d(D))
        return J.reshape(-1, 2), D.reshape(-1, 2)

    def indep_expectation(self, tree: Tree, rates2):
        J, D = np.zeros(tree.n_nodes * 2), np.zeros(tree.n_nodes * 2)
        self._indep_expectation(tree.n_nodes, _u64(tree.subtree_sizes), _u64(tree.parent_ids),
                                _d(tree.branches), _d(np.ascontiguousarray(rates2, np.float64)), _d(J), _d(D))
        return J.reshape(-1, 2), D.reshape(-1, 2)

    def indep_update_paths(self, tree: Tree, rates2, seed: int) -> None:
        self._indep_update_paths(tree.n_nodes, _u64(tree.subtree_sizes), _u64(tree.parent_ids),
                                 _d(tree.branches), _d(np.ascontiguousarray(rates2, np.float64)), seed)

    def indep_estimate(self, tree: Tree, J, D, rates2, with_branches: bool):
        r = np.ascontiguousarray(rates2, np.float64).copy()
        br = np.ascontiguousarray(tree.branches, np.float64).copy()
        self._indep_estimate(tree.n_nodes, _u64(tree.subtree_sizes), _u64(tree.parent_ids), _d(br),
                             _d(np.ascontiguousarray(J, np.float64).reshape(-1).copy()),
                             _d(np.ascontiguousarray(D, np.float64).reshape(-1).copy()), _d(r),
                             int(with_branches))
        return r, br

    def read_paths_file(self, filename: str, timing: bool = False):
        """The reference's read_paths on a local-paths file (reference build only)."""
        ns, nn = np.zeros(1, np.uint64), np.zeros(1, np.uint64)
        sec = self._read_paths_file(filename.encode(), _u64(ns), _u64(nn))
        if sec < 0:
            raise RuntimeError("cannot read: " + filename)
        self._n_sites, self._n_nodes = int(ns[0]), int(nn[0])
        ps = self.store_paths()
        return (ps, sec) if timing else ps

    def write_paths_file(self, filename: str, names) -> float:
        """The reference's writer statements on the loaded paths; returns seconds."""
        arr = (C.c_char_p * len(names))(*[n.encode() for n in names])
        sec = self._write_paths_file(filename.encode(), len(names), arr)
        if sec < 0:
            raise RuntimeError("bad output file: " + filename)
        return sec

    def get_suff_stats(self) -> Tuple[np.ndarray, np.ndarray]:
        J = np.zeros(self._n_nodes * 8)
        D = np.zeros(self._n_nodes * 8)
        self._get_suff_stats(_d(J), _d(D))
        return J.reshape(-1, 8), D.reshape(-1, 8)

    def tri_llh(self, rates, T) -> np.ndarray:
        out = np.zeros(self._n_sites)
        self._tri_llh(_d(np.ascontiguousarray(rates)), _d(np.ascontiguousarray(T)), _d(out))
        return out

    def site_debug(self, tree: Tree, rates, T, site: int, cap: int = 4096):
        seg_off = np.zeros(tree.n_nodes + 1, np.uint64)
        r0, r1, ln = np.zeros(cap), np.zeros(cap), np.zeros(cap)
        t0, t1 = np.zeros(cap, np.uint64), np.zeros(cap, np.uint64)
        p = np.zeros(2 * cap)
        q = np.zeros(2 * tree.n_nodes)
        root_p0, logp = C.c_double(), C.c_double()
        k = self._site_debug(tree.n_nodes, _u64(tree.subtree_sizes), _u64(tree.parent_ids),
                             _d(tree.branches), _d(np.ascontiguousarray(rates)),
                             _d(np.ascontiguousarray(T)), site, cap, _u64(seg_off),
                             _d(r0), _d(r1), _u64(t0), _u64(t1), _d(ln), _d(p), _d(q),
                             C.byref(root_p0), C.byref(logp))
        assert k != 2 ** 64 - 1, "site_debug capacity exceeded"
        k = int(k)
        return dict(seg_off=seg_off, rate0=r0[:k], rate1=r1[:k], trip0=t0[:k],
                    trip1=t1[:k], len=ln[:k], p=p[:2 * k].reshape(-1, 2),
                    q=q.reshape(-1, 2), root_p0=root_p0.value, proposal_logp=logp.value)

    # -- sampler ------------------------------------------------------------------
    def seed(self, seed: int) -> None:
        self._seed(seed)

    def run_mcmc(self, tree: Tree, rates, T, burn_in: int, batch: int, reset: bool = True):
        """Returns (J[n_nodes,8], D[n_nodes,8], acc_rate, seconds)."""
        J = np.zeros(tree.n_nodes * 8)
        D = np.zeros(tree.n_nodes * 8)
        acc, sec = C.c_double(), C.c_double()
        self._keep = (np.ascontiguousarray(rates, np.float64).copy(),
                      np.ascontiguousarray(T, np.float64).copy(),
                      tree.subtree_sizes.copy(), tree.parent_ids.copy(),
                      tree.branches.copy())
        r, t, ss, pi, br = self._keep
        self._run_mcmc(tree.n_nodes, _u64(ss), _u64(pi), _d(br), _d(r), _d(t),
                       burn_in, batch, int(reset), _d(J), _d(D), C.byref(acc), C.byref(sec))
        return J.reshape(-1, 8), D.reshape(-1, 8), acc.value, sec.value

    def time_suff_stats(self, reps: int) -> float:
        J = np.zeros(self._n_nodes * 8)
        D = np.zeros(self._n_nodes * 8)
        return self._time_suff_stats(reps, _d(J), _d(D))

    def end_cond_sample(self, r0, r1, a, b, T, cap: int = 256) -> np.ndarray:
        out = np.zeros(cap)
        n = int(self._end_cond_sample(r0, r1, a, b, T, cap, _d(out)))
        return out[:min(n, cap)]

    # -- model / M-step -----------------------------------------------------------
    def model_from_params(self, T00, T11, b00, b11, scale=True):
        rates, T, Q, bl = np.zeros(8), np.zeros(4), np.zeros(4), np.zeros(4)
        self._model_from_params(T00, T11, b00, b11, int(scale), _d(rates), _d(T), _d(Q), _d(bl))
        return rates, T, Q, bl

    def model_rebuild(self, rates):
        T, Q, bl = np.zeros(4), np.zeros(4), np.zeros(4)
        self._model_rebuild(_d(np.ascontiguousarray(rates, np.float64)), _d(T), _d(Q), _d(bl))
        return T, Q, bl

    def rate_scaling_factor(self, rates) -> float:
        return self._rate_scaling_factor(_d(np.ascontiguousarray(rates, np.float64)))

    def estimate_rates(self, J, D, rates, tol=1e-10):
        J = np.ascontiguousarray(J, np.float64).reshape(-1)
        D = np.ascontiguousarray(D, np.float64).reshape(-1)
        r = np.array(rates, np.float64)
        T, bl = np.zeros(4), np.zeros(4)
        llh = self._estimate_rates(len(J) // 8, _d(J), _d(D), tol, _d(r), _d(T), _d(bl))
        return llh, r, T, bl

    def estimate_rates_and_branches(self, J, D, rates, branches, tol=1e-10):
        J = np.ascontiguousarray(J, np.float64).reshape(-1)
        D = np.ascontiguousarray(D, np.float64).reshape(-1)
        r = np.array(rates, np.float64)
        br = np.array(branches, np.float64)
        T, bl = np.zeros(4), np.zeros(4)
        llh = self._estimate_rates_and_branches(len(J) // 8, _d(J), _d(D), tol, _d(r),
                                                _d(br), _d(T), _d(bl))
        return llh, r, br, T, bl

    def set_one_change(self, rates, branches):
        r = np.array(rates, np.float64)
        br = np.array(branches, np.float64)
        self._set_one_change(_d(r), len(br), _d(br))
        return r, br

    def scale_jump_times(self, new_branches) -> None:
        br = np.ascontiguousarray(new_branches, np.float64)
        self._scale_jump_times(len(br), _d(br))

    def log_likelihood(self, J, D, rates) -> float:
        return self._log_likelihood(_d(np.ascontiguousarray(J, np.float64)),
                                    _d(np.ascontiguousarray(D, np.float64)),
                                    _d(np.ascontiguousarray(rates, np.float64)))

    def gradient(self, J, D, rates) -> np.ndarray:
        g = np.zeros(8)
        self._gradient(_d(np.ascontiguousarray(J, np.float64)),
                       _d(np.ascontiguousarray(D, np.float64)),
                       _d(np.ascontiguousarray(rates, np.float64)), _d(g))
        return g


def build_oracle() -> None:
    subprocess.run(["make", "-s", "-C", ORACLE_DIR, "oracle", "ref"], check=True)


_oracle: Optional[CpuLib] = None
_ref: Optional[CpuLib] = None


def oracle() -> CpuLib:
    global _oracle
    if _oracle is None:
        p = os.path.join(ORACLE_DIR, "liboracle.so")
        if not os.path.exists(p):
            build_oracle()
        _oracle = CpuLib(p, "orc_")
    return _oracle


def ref_available() -> bool:
    return os.path.exists(os.path.join(ORACLE_DIR, "_ref", "libepievo_ref.so"))


def ref() -> CpuLib:
    global _ref
    if _ref is None:
        _ref = CpuLib(os.path.join(ORACLE_DIR, "_ref", "libepievo_ref.so"), "ref_")
    return _ref


def best_cpu() -> CpuLib:
    """The real reference when its build is present, else the C port."""
    return ref() if ref_available() else oracle()
