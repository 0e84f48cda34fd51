"""Thin object wrapper over the C-ABI context (one context = one GPU)."""
from __future__ import annotations

import ctypes as C
from typing import Optional, Tuple

import numpy as np

from . import _lib as L
from .paths import PathSet, Tree


def _d(a):
    return a.ctypes.data_as(L.dp)


def _u64(a):
    return a.ctypes.data_as(L.u64p)


def _u8(a):
    return a.ctypes.data_as(L.u8p)


def _f64(x) -> np.ndarray:
    return np.ascontiguousarray(x, dtype=np.float64)


class EpvError(RuntimeError):
    pass


class Context:
    def __init__(self, device: int = 0):
        self.lib = L.load()
        self.h = L.ctx_p()
        if self.lib.epv_create(C.byref(self.h), device) != 0:
            self.h = None
            raise EpvError("epv_create failed: no usable CUDA device %d (no CPU fallback)" % device)
        self.n_nodes = 0
        self.n_sites = 0

    def close(self):
        if getattr(self, "h", None):
            self.lib.epv_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc: int):
        if rc != 0:
            raise EpvError(self.lib.epv_last_error(self.h).decode())

    # -- setup --------------------------------------------------------------------
    def set_stream(self, cuda_stream_ptr: int):
        self._ck(self.lib.epv_set_stream(self.h, C.c_void_p(cuda_stream_ptr)))

    def synchronize(self):
        self._ck(self.lib.epv_synchronize(self.h))

    def set_tree(self, tree: Tree):
        ss = np.ascontiguousarray(tree.subtree_sizes, np.uint64)
        pi = np.ascontiguousarray(tree.parent_ids, np.uint64)
        br = _f64(tree.branches)
        self._ck(self.lib.epv_set_tree(self.h, tree.n_nodes, _u64(ss), _u64(pi), _d(br)))
        self.n_nodes = tree.n_nodes

    def set_model(self, rates, T):
        self._ck(self.lib.epv_set_model(self.h, _d(_f64(rates)), _d(_f64(T))))

    def set_shard(self, n_global, site_begin, site_end, halo_lo, halo_hi):
        self._ck(self.lib.epv_set_shard(self.h, n_global, site_begin, site_end, halo_lo, halo_hi))

    def upload_paths(self, ps: PathSet):
        j = ps.jumps if len(ps.jumps) else np.zeros(1)
        self._ck(self.lib.epv_upload_paths(self.h, ps.n_sites, ps.n_nodes, _u8(ps.init_state),
                                           _u64(ps.row_off), _d(j), _d(ps.tot_time)))
        self.n_sites = ps.n_sites

    def upload_raw(self, n_sites, n_nodes, init_ptr, off_ptr, jumps_ptr, tot_time):
        """Upload from raw host addresses (e.g. pinned torch tensors)."""
        tt = _f64(tot_time)
        self._ck(self.lib.epv_upload_paths(
            self.h, n_sites, n_nodes, C.cast(init_ptr, L.u8p), C.cast(off_ptr, L.u64p),
            C.cast(jumps_ptr, L.dp), _d(tt)))
        self.n_sites = n_sites

    def paths_size(self) -> Tuple[int, int, int]:
        a, b, c = C.c_uint64(), C.c_uint64(), C.c_uint64()
        self._ck(self.lib.epv_paths_size(self.h, C.byref(a), C.byref(b), C.byref(c)))
        return a.value, b.value, c.value

    def download_paths(self, node_names=()) -> PathSet:
        n_sites, n_nodes, n_jumps = self.paths_size()
        init = np.zeros(n_sites * n_nodes, np.uint8)
        off = np.zeros(n_sites * n_nodes + 1, np.uint64)
        jumps = np.zeros(max(1, n_jumps))
        tot = np.zeros(n_nodes)
        self._ck(self.lib.epv_download_paths(self.h, _u8(init), _u64(off), _d(jumps), _d(tot)))
        return PathSet(n_sites, n_nodes, init, off, jumps[:n_jumps].copy(), tot, list(node_names))

    def download_raw(self, init_ptr, off_ptr, jumps_ptr):
        tot = np.zeros(self.n_nodes)
        self._ck(self.lib.epv_download_paths(
            self.h, C.cast(init_ptr, L.u8p), C.cast(off_ptr, L.u64p), C.cast(jumps_ptr, L.dp),
            _d(tot)))

    # compact exchange format: u16 row codes (count | init << 15) + jump times
    def upload_paths_compact(self, ps: PathSet):
        j = ps.jumps if len(ps.jumps) else np.zeros(1)
        code = ps.row_code()
        self._ck(self.lib.epv_upload_paths_compact(
            self.h, ps.n_sites, ps.n_nodes, code.ctypes.data_as(L.u16p), _d(j), len(ps.jumps),
            _d(ps.tot_time)))
        self.n_sites = ps.n_sites

    def download_paths_compact(self, node_names=()) -> PathSet:
        n_sites, n_nodes, n_jumps = self.paths_size()
        code = np.zeros(n_sites * n_nodes, np.uint16)
        jumps = np.zeros(max(1, n_jumps))
        tot = np.zeros(n_nodes)
        n = C.c_uint64()
        self._ck(self.lib.epv_download_paths_compact(
            self.h, code.ctypes.data_as(L.u16p), _d(jumps), len(jumps), C.byref(n), _d(tot)))
        return PathSet.from_row_code(n_sites, n_nodes, code, jumps[:n.value].copy(), tot,
                                     list(node_names))

    def upload_compact_raw(self, n_sites, n_nodes, code_ptr, jumps_ptr, n_jumps, tot_time):
        tt = _f64(tot_time)
        self._ck(self.lib.epv_upload_paths_compact(
            self.h, n_sites, n_nodes, C.cast(code_ptr, L.u16p), C.cast(jumps_ptr, L.dp), n_jumps,
            _d(tt)))
        self.n_sites = n_sites

    def download_compact_raw(self, code_ptr, jumps_ptr, jumps_cap) -> int:
        tot = np.zeros(self.n_nodes)
        n = C.c_uint64()
        self._ck(self.lib.epv_download_paths_compact(
            self.h, C.cast(code_ptr, L.u16p), C.cast(jumps_ptr, L.dp), jumps_cap, C.byref(n),
            _d(tot)))
        return n.value

    def jump_counts(self) -> np.ndarray:
        out = np.zeros(self.n_nodes, np.uint64)
        self._ck(self.lib.epv_jump_counts(self.h, _u64(out)))
        return out

    # -- sampler ------------------------------------------------------------------
    def reset(self):
        self._ck(self.lib.epv_reset(self.h))

    def sweep(self, seed: int, sweep_idx: int) -> int:
        n = C.c_uint64()
        self._ck(self.lib.epv_sweep(self.h, seed, sweep_idx, C.byref(n)))
        return n.value

    def colour_pass(self, colour: int, seed: int, sweep_idx: int):
        self._ck(self.lib.epv_pass(self.h, colour, seed, sweep_idx))

    def accept_count(self, reset: bool = True) -> int:
        n = C.c_uint64()
        self._ck(self.lib.epv_accept_count(self.h, int(reset), C.byref(n)))
        return n.value

    def run_mcmc(self, burn_in: int, batch: int, seed: int, first_sweep: int = 0):
        J = np.zeros(self.n_nodes * 8)
        D = np.zeros(self.n_nodes * 8)
        acc = C.c_double()
        self._ck(self.lib.epv_run_mcmc(self.h, burn_in, batch, seed, first_sweep, _d(J), _d(D),
                                       C.byref(acc)))
        return J.reshape(-1, 8), D.reshape(-1, 8), acc.value

    def run_mcmc_raw(self, burn_in: int, batch: int, seed: int, first_sweep: int = 0):
        J = np.zeros(self.n_nodes * 8)
        D = np.zeros(self.n_nodes * 8)
        n = C.c_uint64()
        self._ck(self.lib.epv_run_mcmc_raw(self.h, burn_in, batch, seed, first_sweep, _d(J),
                                           _d(D), C.byref(n)))
        return J.reshape(-1, 8), D.reshape(-1, 8), n.value

    def mcem_iteration(self, burn_in, batch, seed, first_sweep, rates, T, branches,
                       optimize_branches=False, param_tol=1e-10):
        """One device-resident MCEM iteration; returns the updated (rates, T, branches,
        baseline, acc_rate, llh)."""
        r, t, br, bl = _f64(rates).copy(), _f64(T).copy(), _f64(branches).copy(), np.zeros(4)
        acc, llh = C.c_double(), C.c_double()
        self._ck(self.lib.epv_mcem_iteration(self.h, burn_in, batch, seed, first_sweep,
                                             int(optimize_branches), param_tol, _d(r), _d(t), _d(br),
                                             _d(bl), C.byref(acc), C.byref(llh)))
        return r, t, br, bl, acc.value, llh.value

    # -- multi-GPU halo exchange (device buffers given as raw addresses) ------------------
    def halo_record_bytes(self) -> int:
        n = C.c_uint64()
        self._ck(self.lib.epv_halo_record_bytes(self.h, C.byref(n)))
        return n.value

    def halo_pack(self, colour: int, rec_lo_ptr: Optional[int], rec_hi_ptr: Optional[int]):
        self._ck(self.lib.epv_halo_pack(self.h, colour, C.c_void_p(rec_lo_ptr), C.c_void_p(rec_hi_ptr)))

    def halo_unpack(self, colour: int, rec_lo_ptr: Optional[int], rec_hi_ptr: Optional[int]):
        self._ck(self.lib.epv_halo_unpack(self.h, colour, C.c_void_p(rec_lo_ptr), C.c_void_p(rec_hi_ptr)))

    def halo_capacity(self) -> Tuple[int, int]:
        """(jump capacity of a halo row, capacity the packed boundary rows ask for)."""
        cap, need = C.c_uint64(), C.c_uint64()
        self._ck(self.lib.epv_halo_capacity(self.h, C.byref(cap), C.byref(need)))
        return cap.value, need.value

    def halo_set_capacity(self, capacity: int):
        self._ck(self.lib.epv_halo_set_capacity(self.h, capacity))

    def halo_need_async(self, dst_device_ptr: int, cuda_stream_ptr: Optional[int] = None):
        """Longest boundary row packed so far -> one uint64 at dst_device_ptr, enqueued (no synchronisation)."""
        self._ck(self.lib.epv_halo_need_async(self.h, C.c_void_p(dst_device_ptr), C.c_void_p(cuda_stream_ptr)))

    def set_side_stream(self, cuda_stream_ptr: int):
        self._ck(self.lib.epv_set_side_stream(self.h, C.c_void_p(cuda_stream_ptr)))

    def pass_begin(self, colour: int, seed: int, sweep_idx: int, rec_lo_ptr: Optional[int], rec_hi_ptr: Optional[int]):
        self._ck(self.lib.epv_pass_begin(self.h, colour, seed, sweep_idx, C.c_void_p(rec_lo_ptr), C.c_void_p(rec_hi_ptr)))

    def pass_end(self, colour: int, rec_lo_ptr: Optional[int], rec_hi_ptr: Optional[int]):
        self._ck(self.lib.epv_pass_end(self.h, colour, C.c_void_p(rec_lo_ptr), C.c_void_p(rec_hi_ptr)))

    # -- statistics -----------------------------------------------------------------
    def stats_words(self) -> int:
        n = C.c_uint64()
        self._ck(self.lib.epv_stats_words(self.h, C.byref(n)))
        return n.value

    def stats_begin(self):
        self._ck(self.lib.epv_stats_begin(self.h))

    def stats_accumulate(self):
        self._ck(self.lib.epv_stats_accumulate(self.h))

    def stats_export(self, dev_ptr: int):
        self._ck(self.lib.epv_stats_export(self.h, C.c_void_p(dev_ptr)))

    def stats_fold(self, dev_ptr: Optional[int] = None):
        J, D = np.zeros(self.n_nodes * 8), np.zeros(self.n_nodes * 8)
        self._ck(self.lib.epv_stats_fold(self.h, C.c_void_p(dev_ptr), _d(J), _d(D)))
        return J.reshape(-1, 8), D.reshape(-1, 8)

    def suffstats_static(self):
        """epv_suffstats_static: the J/D pass with its scratch in the library's static device array
        (no allocation; trees of at most 64 nodes)."""
        J = np.zeros(self.n_nodes * 8)
        D = np.zeros(self.n_nodes * 8)
        self._ck(self.lib.epv_suffstats_static(self.h, _d(J), _d(D)))
        return J.reshape(-1, 8), D.reshape(-1, 8)

    def suffstats(self):
        J = np.zeros(self.n_nodes * 8)
        D = np.zeros(self.n_nodes * 8)
        self._ck(self.lib.epv_suffstats(self.h, _d(J), _d(D)))
        return J.reshape(-1, 8), D.reshape(-1, 8)

    def llh_grad(self, rates):
        llh = C.c_double()
        g = np.zeros(8)
        self._ck(self.lib.epv_llh_grad(self.h, _d(_f64(rates)), C.byref(llh), _d(g)))
        return llh.value, g

    def scale_jump_times(self, new_branches):
        self._ck(self.lib.epv_scale_jump_times(self.h, _d(_f64(new_branches))))

    # -- site-independent E-step (epievo_initialization) ---------------------------------
    def indep_suffstats(self):
        """Sums over owned sites of the per-state J, D of every branch, [n_nodes, 2]."""
        J, D = np.zeros(self.n_nodes * 2), np.zeros(self.n_nodes * 2)
        self._ck(self.lib.epv_indep_suffstats(self.h, _d(J), _d(D)))
        return J.reshape(-1, 2), D.reshape(-1, 2)

    def indep_expectation(self, rates2):
        J, D = np.zeros(self.n_nodes * 2), np.zeros(self.n_nodes * 2)
        self._ck(self.lib.epv_indep_expectation(self.h, _d(_f64(rates2)), _d(J), _d(D)))
        return J.reshape(-1, 2), D.reshape(-1, 2)

    def indep_update_paths(self, rates2, seed: int, iteration: int = 0):
        self._ck(self.lib.epv_indep_update_paths(self.h, _d(_f64(rates2)), seed, iteration))

    def simulate_adopt(self, n_sites: int, seed: int, root_states=None) -> int:
        """epv_simulate + epv_simulate_adopt: the simulated histories become the context's paths
        without leaving the device.  Returns the number of jumps."""
        nj = C.c_uint64()
        root = None if root_states is None else _u8(np.ascontiguousarray(root_states, np.uint8))
        self._ck(self.lib.epv_simulate(self.h, n_sites, root, seed, C.byref(nj)))
        self._ck(self.lib.epv_simulate_adopt(self.h))
        self.n_sites = n_sites
        return nj.value

    # -- forward simulation (epievo_sim -> global_jumps_to_paths) --------------------------
    def simulate(self, n_sites: int, seed: int, tot_time, root_states=None, node_names=()) -> PathSet:
        """Histories of n_sites evolved down the context's tree under its model; root drawn from
        the horizontal chain unless given.  tot_time: the tree's branch lengths [n_nodes]."""
        nj = C.c_uint64()
        root = None if root_states is None else _u8(np.ascontiguousarray(root_states, np.uint8))
        self._ck(self.lib.epv_simulate(self.h, n_sites, root, seed, C.byref(nj)))
        init = np.zeros(n_sites * self.n_nodes, np.uint8)
        off = np.zeros(n_sites * self.n_nodes + 1, np.uint64)
        jumps = np.zeros(max(1, nj.value))
        self._ck(self.lib.epv_simulate_fetch(self.h, _u8(init), _u64(off), _d(jumps)))
        tt = _f64(tot_time).copy()
        tt[0] = 0.0
        return PathSet(n_sites, self.n_nodes, init, off, jumps[:nj.value].copy(), tt, list(node_names))

    # -- parity probes ----------------------------------------------------------------
    def site_debug(self, site: int, cap: int = 4096):
        nn = self.n_nodes
        k = C.c_uint64()
        seg_off = np.zeros(nn + 1, np.uint64)
        r0, r1, ln = np.zeros(cap), np.zeros(cap), np.zeros(cap)
        t0, t1 = np.zeros(cap, np.uint64), np.zeros(cap, np.uint64)
        p, q = np.zeros(2 * cap), np.zeros(2 * nn)
        rp, lp = C.c_double(), C.c_double()
        self._ck(self.lib.epv_site_debug(self.h, site, cap, C.byref(k), _u64(seg_off), _d(r0),
                                         _d(r1), _u64(t0), _u64(t1), _d(ln), _d(p), _d(q),
                                         C.byref(rp), C.byref(lp)))
        n = k.value
        return dict(seg_off=seg_off, rate0=r0[:n], rate1=r1[:n], trip0=t0[:n], trip1=t1[:n],
                    len=ln[:n], p=p[:2 * n].reshape(-1, 2), q=q.reshape(-1, 2),
                    root_p0=rp.value, proposal_logp=lp.value)

    def tri_llh(self) -> np.ndarray:
        out = np.zeros(self.n_sites)
        self._ck(self.lib.epv_get_tri_llh(self.h, _d(out)))
        return out

    def end_cond_sample(self, r0, r1, a, b, T, seed, n, cap=32):
        nj = np.zeros(n, np.uint32)
        j = np.zeros(n * cap)
        self._ck(self.lib.epv_end_cond_sample(self.h, r0, r1, a, b, T, seed, n, cap,
                                              nj.ctypes.data_as(L.u32p), _d(j)))
        return nj, j.reshape(n, cap)

    # -- instrumentation ----------------------------------------------------------------
    def launch_count(self) -> int:
        return int(self.lib.epv_launch_count(self.h))

    def recovery_count(self) -> int:
        return int(self.lib.epv_recovery_count(self.h))

    def enable_kernel_timing(self, on: bool):
        self._ck(self.lib.epv_enable_kernel_timing(self.h, int(on)))

    def mh_kernel_time(self, reset: bool = True):
        ms, n = C.c_double(), C.c_uint64()
        self._ck(self.lib.epv_mh_kernel_time(self.h, int(reset), C.byref(ms), C.byref(n)))
        return ms.value, n.value

    def rebuild_kernel_time(self) -> float:
        ms = C.c_double()
        self._ck(self.lib.epv_rebuild_kernel_time(self.h, C.byref(ms)))
        return ms.value


class MultiContext:
    """Several GPUs behind one handle in one process (epv_multi_*, csrc/api_multi.cu): the calls of
    Context with the same meaning, over contiguous site blocks with peer-copied halo records."""

    def __init__(self, n_shards: int, devices=None):
        self.lib = L.load()
        self.h = L.ctx_p()
        dev = None
        if devices is not None:
            assert len(devices) == n_shards
            dev = (C.c_int * n_shards)(*devices)
        if self.lib.epv_multi_create(C.byref(self.h), n_shards, dev) != 0:
            self.h = None
            raise EpvError("epv_multi_create failed (no usable CUDA devices; there is no CPU fallback)")
        self.n_shards = n_shards
        self.n_nodes = 0

    def close(self):
        if getattr(self, "h", None):
            self.lib.epv_multi_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc: int):
        if rc != 0:
            raise EpvError(self.lib.epv_multi_last_error(self.h).decode())

    def set_tree(self, tree: Tree):
        ss = np.ascontiguousarray(tree.subtree_sizes, np.uint64)
        pi = np.ascontiguousarray(tree.parent_ids, np.uint64)
        self._ck(self.lib.epv_multi_set_tree(self.h, tree.n_nodes, _u64(ss), _u64(pi), _d(_f64(tree.branches))))
        self.n_nodes = tree.n_nodes

    def set_model(self, rates, T):
        self._ck(self.lib.epv_multi_set_model(self.h, _d(_f64(rates)), _d(_f64(T))))

    def upload_paths(self, ps: PathSet):
        j = ps.jumps if len(ps.jumps) else np.zeros(1)
        self._ck(self.lib.epv_multi_upload_paths(self.h, ps.n_sites, ps.n_nodes, _u8(ps.init_state),
                                                 _u64(ps.row_off), _d(j), _d(ps.tot_time)))

    def upload_raw(self, n_sites, n_nodes, init_ptr, off_ptr, jumps_ptr, tot_time):
        self._ck(self.lib.epv_multi_upload_paths(
            self.h, n_sites, n_nodes, C.cast(init_ptr, L.u8p), C.cast(off_ptr, L.u64p),
            C.cast(jumps_ptr, L.dp), _d(_f64(tot_time))))

    def paths_size(self) -> Tuple[int, int, int]:
        a, b, c = C.c_uint64(), C.c_uint64(), C.c_uint64()
        self._ck(self.lib.epv_multi_paths_size(self.h, C.byref(a), C.byref(b), C.byref(c)))
        return a.value, b.value, c.value

    def download_paths(self, node_names=()) -> PathSet:
        n_sites, n_nodes, n_jumps = self.paths_size()
        init = np.zeros(n_sites * n_nodes, np.uint8)
        off = np.zeros(n_sites * n_nodes + 1, np.uint64)
        jumps = np.zeros(max(1, n_jumps))
        tot = np.zeros(n_nodes)
        self._ck(self.lib.epv_multi_download_paths(self.h, _u8(init), _u64(off), _d(jumps), _d(tot)))
        return PathSet(n_sites, n_nodes, init, off, jumps[:n_jumps].copy(), tot, list(node_names))

    def reset(self):
        self._ck(self.lib.epv_multi_reset(self.h))

    def sweep(self, seed: int, sweep_idx: int) -> int:
        n = C.c_uint64()
        self._ck(self.lib.epv_multi_sweep(self.h, seed, sweep_idx, C.byref(n)))
        return n.value

    def run_mcmc(self, burn_in: int, batch: int, seed: int, first_sweep: int = 0):
        J, D, acc = np.zeros(self.n_nodes * 8), np.zeros(self.n_nodes * 8), C.c_double()
        self._ck(self.lib.epv_multi_run_mcmc(self.h, burn_in, batch, seed, first_sweep, _d(J), _d(D), C.byref(acc)))
        return J.reshape(-1, 8), D.reshape(-1, 8), acc.value

    def suffstats(self):
        J, D = np.zeros(self.n_nodes * 8), np.zeros(self.n_nodes * 8)
        self._ck(self.lib.epv_multi_suffstats(self.h, _d(J), _d(D)))
        return J.reshape(-1, 8), D.reshape(-1, 8)

    def scale_jump_times(self, new_branches):
        self._ck(self.lib.epv_multi_scale_jump_times(self.h, _d(_f64(new_branches))))

    def mcem_iteration(self, burn_in, batch, seed, first_sweep, rates, T, branches,
                       optimize_branches=False, param_tol=1e-10):
        r, t, br, bl = _f64(rates).copy(), _f64(T).copy(), _f64(branches).copy(), np.zeros(4)
        acc, llh = C.c_double(), C.c_double()
        self._ck(self.lib.epv_multi_mcem_iteration(self.h, burn_in, batch, seed, first_sweep,
                                                   int(optimize_branches), param_tol, _d(r), _d(t), _d(br),
                                                   _d(bl), C.byref(acc), C.byref(llh)))
        return r, t, br, bl, acc.value, llh.value

    def launch_count(self) -> int:
        return int(self.lib.epv_multi_launch_count(self.h))
