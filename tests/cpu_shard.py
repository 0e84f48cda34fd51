"""CPU stand-in for one GPU shard (TEST INFRASTRUCTURE): the same interface
epievo_b200.sharded drives (ctx.colour_pass / suffstats / accept_count, pack / unpack,
send_* / recv_* byte tensors), implemented over the oracle's colour-class sweeps with
per-site reseeded mt19937.  Lets the sharding PROTOCOL -- which site travels after which
pass, who owns which tri_llh entry, the J/D all-reduce -- run under torch.distributed
with the gloo backend on machines without a GPU.  One shard per process (the oracle
keeps its paths in process-global state)."""
from __future__ import annotations

import ctypes as C
import pickle

import numpy as np
import torch

import cpu_libs
from epievo_b200 import sharded as S

REC_BYTES = 1 << 16


class _CpuCtx:
    def __init__(self, lib, tree, rates, T, n_global, lo, hi, halo_lo, halo_hi):
        self.o, self.tree, self.n_nodes = lib, tree, tree.n_nodes
        self.rates, self.T = np.ascontiguousarray(rates), np.ascontiguousarray(T)
        self.n_global, self.lo, self.hi, self.halo_lo, self.halo_hi = n_global, lo, hi, halo_lo, halo_hi
        self.acc = 0
        L = lib.lib
        L.orc_sampler_init.argtypes = [C.c_uint64] + [C.c_void_p] * 5
        L.orc_colour_pass.restype = C.c_uint64
        L.orc_colour_pass.argtypes = [C.c_int, C.c_uint64, C.c_uint64, C.c_int64, C.c_uint64, C.c_uint64, C.c_uint64]
        L.orc_get_tri_cache.argtypes = [C.c_void_p]
        L.orc_set_tri_cache.argtypes = [C.c_uint64, C.c_double]
        L.orc_set_row.argtypes = [C.c_uint64, C.c_uint64, C.c_int, C.c_uint64, C.c_void_p]

    def upload_paths(self, ps):
        self.n_sites = ps.n_sites
        self.o.load_paths(ps)
        t = self.tree
        self._keep = (t.subtree_sizes.copy(), t.parent_ids.copy(), np.array(t.branches, np.float64))
        self.o.lib.orc_sampler_init(self.n_nodes, *[a.ctypes.data for a in self._keep],
                                    self.rates.ctypes.data, self.T.ctypes.data)

    def colour_pass(self, colour, seed, sweep):
        self.acc += self.o.lib.orc_colour_pass(colour, seed, sweep, self.lo - self.halo_lo, self.n_global,
                                               self.halo_lo, self.n_sites - self.halo_hi)

    def accept_count(self, reset=True):
        n = self.acc
        if reset:
            self.acc = 0
        return n

    def suffstats(self):
        """J/D over the owned, globally interior sites (what epv_suffstats counts)."""
        ps = self.o.store_paths()
        g0 = self.lo - self.halo_lo
        a = max(self.halo_lo, 1 - g0)
        b = min(self.n_sites - self.halo_hi, self.n_global - 1 - g0)
        J, D = np.zeros((self.n_nodes, 8)), np.zeros((self.n_nodes, 8))
        for v in range(1, self.n_nodes):
            for s in range(a, b):
                j, d = self.o.add_suff_stats(*[(ps.init(v, x), ps.row(v, x)) for x in (s - 1, s, s + 1)],
                                             float(ps.tot_time[v]))
                J[v] += j
                D[v] += d
        return J, D

    def tri_llh(self):
        out = np.zeros(self.n_sites)
        self.o.lib.orc_get_tri_cache(out.ctypes.data)
        return out

    def download_paths(self):
        return self.o.store_paths()

    # the stand-in's records are pickles of any size: nothing to agree on
    def halo_capacity(self):
        return 48, 0

    def halo_set_capacity(self, capacity):
        pass


class CpuShard:
    def __init__(self, tree, rates, T, n_global, world, rank):
        self.rank, self.world, self.n_global = rank, world, n_global
        self.lo, self.hi, self.halo_lo, self.halo_hi = S.shard_bounds(n_global, world, rank)
        self.ctx = _CpuCtx(cpu_libs.oracle(), tree, rates, T, n_global, self.lo, self.hi,
                           self.halo_lo, self.halo_hi)
        mk = lambda: torch.zeros(REC_BYTES, dtype=torch.uint8)
        self.send_lo, self.send_hi, self.recv_lo, self.recv_hi = mk(), mk(), mk(), mk()
        self.side, self.device = None, torch.device("cpu")

    def ensure_records(self):
        pass

    # epv_pass_begin / epv_pass_end: update + pack, then unpack (no streams to overlap on the CPU)
    def begin(self, colour, seed, sweep):
        self.ctx.colour_pass(colour, seed, sweep)
        self.pack(colour)

    def end(self, colour):
        self.unpack(colour)

    # the same site arithmetic as csrc/api_halo.cu
    def _pack_sites(self, colour):
        own_lo, own_hi = self.halo_lo, self.ctx.n_sites - self.halo_hi
        lo = own_lo + colour if self.halo_lo else None
        hi = own_hi - 3 + (colour - (own_hi - 3)) % 3 if self.halo_hi else None
        return lo, hi

    def _unpack_sites(self, colour):
        own_hi = self.ctx.n_sites - self.halo_hi
        return (colour if self.halo_lo else None), (own_hi + (colour - own_hi) % 3 if self.halo_hi else None)

    def pack(self, colour):
        ps, tri = self.ctx.download_paths(), self.ctx.tri_llh()
        for site, buf in zip(self._pack_sites(colour), (self.send_lo, self.send_hi)):
            if site is None:
                continue
            rec = dict(g=self.lo - self.halo_lo + site, tri=tri[site - 1:site + 2].copy(),
                       init=[ps.init(v, site) for v in range(ps.n_nodes)],
                       rows=[ps.row(v, site).copy() for v in range(ps.n_nodes)])
            raw = pickle.dumps(rec)
            assert len(raw) + 8 <= REC_BYTES
            buf[:8] = torch.from_numpy(np.frombuffer(np.int64(len(raw)).tobytes(), np.uint8).copy())
            buf[8:8 + len(raw)] = torch.from_numpy(np.frombuffer(raw, np.uint8).copy())

    def unpack(self, colour):
        L = self.ctx.o.lib
        for site, buf in zip(self._unpack_sites(colour), (self.recv_lo, self.recv_hi)):
            if site is None:
                continue
            n = int(np.frombuffer(buf[:8].numpy().tobytes(), np.int64)[0])
            rec = pickle.loads(buf[8:8 + n].numpy().tobytes())
            assert rec["g"] == self.lo - self.halo_lo + site, "halo record for the wrong site"
            for v in range(1, self.ctx.n_nodes):
                j = np.ascontiguousarray(rec["rows"][v], np.float64)
                L.orc_set_row(site, v, int(rec["init"][v]), len(j), j.ctypes.data if len(j) else None)
            for k in range(3):
                if 0 <= site - 1 + k < self.ctx.n_sites:
                    L.orc_set_tri_cache(site - 1 + k, float(rec["tri"][k]))
