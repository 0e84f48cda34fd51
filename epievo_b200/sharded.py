"""Multi-GPU driver: contiguous site blocks, one per GPU, 3-site halos, one tiny
neighbour exchange per colour pass and one all-reduce of the J/D sums per E-step
(SURVEY 8e).  The transport is pluggable so the same protocol code runs

* across processes with torch.distributed (NCCL on device buffers; one rank per GPU), and
* inside one process over several contexts (``LocalExchange``: device-to-device copies),
  which is how the protocol is tested on a single GPU.

Determinism: the Philox streams are keyed by the GLOBAL site index, so a sharded run
reproduces the single-GPU run bit for bit (tests/test_gpu_sharded.py).
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import numpy as np

from .paths import PathSet

HALO = 3


def partition(n_global: int, world: int) -> List[Tuple[int, int]]:
    """Contiguous blocks whose starts are multiples of 3 (colour alignment)."""
    n3 = n_global // 3
    cuts = [3 * ((n3 * r) // world) for r in range(world)] + [n_global]
    blocks = [(cuts[r], cuts[r + 1]) for r in range(world)]
    for lo, hi in blocks:
        if hi - lo < 2 * HALO:
            raise ValueError("too few sites for %d shards" % world)
    return blocks


def shard_bounds(n_global: int, world: int, rank: int) -> Tuple[int, int, int, int]:
    lo, hi = partition(n_global, world)[rank]
    return lo, hi, (HALO if rank > 0 else 0), (HALO if rank + 1 < world else 0)


def local_paths(ps: PathSet, n_global: int, world: int, rank: int) -> PathSet:
    lo, hi, hl, hh = shard_bounds(n_global, world, rank)
    return ps.site_slice(lo - hl, hi + hh)


class Shard:
    """One context + its four device-side halo records + the side stream its halo traffic runs on."""

    def __init__(self, ctx, n_global: int, world: int, rank: int, torch_device, side_stream=None):
        import torch
        self.ctx, self.rank, self.world, self.n_global = ctx, rank, world, n_global
        self.device = torch_device
        self.lo, self.hi, self.halo_lo, self.halo_hi = shard_bounds(n_global, world, rank)
        ctx.set_shard(n_global, self.lo, self.hi, self.halo_lo, self.halo_hi)
        # a torch stream, so that NCCL send/recv (or the test's tensor copies) can be enqueued on it
        self.side = side_stream if side_stream is not None else torch.cuda.Stream(device=torch_device)
        ctx.set_side_stream(self.side.cuda_stream)
        self.rec_bytes = 0
        self.ensure_records()

    def ensure_records(self):
        """(Re)allocates the records for the context's current halo-row capacity."""
        import torch
        nb = self.ctx.halo_record_bytes()
        if nb == self.rec_bytes:
            return
        self.ctx.synchronize()
        mk = lambda: torch.zeros(nb, dtype=torch.uint8, device=self.device)
        self.send_lo, self.send_hi, self.recv_lo, self.recv_hi = mk(), mk(), mk(), mk()
        self.rec_bytes = nb

    def begin(self, colour, seed, sweep):
        self.ctx.pass_begin(colour, seed, sweep, self._p(self.send_lo, self.halo_lo), self._p(self.send_hi, self.halo_hi))

    def end(self, colour):
        self.ctx.pass_end(colour, self._p(self.recv_lo, self.halo_lo), self._p(self.recv_hi, self.halo_hi))

    def _p(self, t, present):
        return t.data_ptr() if present else None

    def pack(self, colour):
        self.ctx.halo_pack(colour, self._p(self.send_lo, self.halo_lo), self._p(self.send_hi, self.halo_hi))

    def unpack(self, colour):
        self.ctx.halo_unpack(colour, self._p(self.recv_lo, self.halo_lo), self._p(self.recv_hi, self.halo_hi))


class DistExchange:
    """torch.distributed transport: rank r talks to r-1 and r+1 only."""

    def __init__(self, shard: Shard):
        self.s = shard

    def exchange(self):
        """ncclSend/ncclRecv pairs with the two ring neighbours, enqueued on the shard's SIDE stream
        (behind the pack of epv_pass_begin, ahead of the unpack of epv_pass_end): nothing here blocks
        the host or the main stream."""
        import torch
        import torch.distributed as dist
        s, ops = self.s, []
        if s.halo_lo:
            ops += [dist.P2POp(dist.isend, s.send_lo, s.rank - 1), dist.P2POp(dist.irecv, s.recv_lo, s.rank - 1)]
        if s.halo_hi:
            ops += [dist.P2POp(dist.isend, s.send_hi, s.rank + 1), dist.P2POp(dist.irecv, s.recv_hi, s.rank + 1)]
        if ops:
            if s.side is None:                      # CPU stand-in shards (gloo tests)
                for w in dist.batch_isend_irecv(ops):
                    w.wait()
                return
            with torch.cuda.stream(s.side):
                for w in dist.batch_isend_irecv(ops):
                    w.wait()

    def allreduce_max(self, value: int) -> int:
        import torch
        import torch.distributed as dist
        t = torch.tensor([value], dtype=torch.int64, device=self.s.device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return int(t.item())

    def allreduce_max_tensor(self, t):
        """Device-side max over ranks, enqueued on the current stream (CapacityWatch)."""
        import torch.distributed as dist
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t

    def allreduce_int(self, t):
        """Exact: int64 sums (the J/D accumulators and the accept count)."""
        import torch.distributed as dist
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return t

    def allreduce_sum(self, arr: np.ndarray) -> np.ndarray:
        import torch
        import torch.distributed as dist
        t = torch.from_numpy(np.ascontiguousarray(arr, np.float64)).to(self.s.send_lo.device)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return t.cpu().numpy()


class LocalExchange:
    """All shards in one process: neighbour records move by device-to-device copy."""

    def __init__(self, shards: Sequence[Shard]):
        self.shards = list(shards)

    def exchange(self):
        """Record copies on the first shard's side stream, ordered after every shard's pack and before
        every shard's unpack."""
        import torch
        main = self.shards[0].side
        for s in self.shards[1:]:
            if s.side is not main:
                main.wait_stream(s.side)
        with torch.cuda.stream(main):
            for a, b in zip(self.shards[:-1], self.shards[1:]):
                b.recv_lo.copy_(a.send_hi)
                a.recv_hi.copy_(b.send_lo)
        for s in self.shards[1:]:
            if s.side is not main:
                s.side.wait_stream(main)

    def allreduce_max(self, value: int) -> int:
        return value

    def allreduce_max_tensor(self, t):
        return t


def colour_pass(shards: Sequence[Shard], transport, colour: int, seed: int, sweep: int, overlap: bool = True):
    """One colour pass over the shards.  overlap: epv_pass_begin / exchange on the side streams /
    epv_pass_end (edge groups first, the records travel while the interior is updated); otherwise the
    serial protocol epv_pass -> pack -> exchange -> unpack on one stream.  Same chain either way."""
    if overlap:
        for s in shards:
            s.begin(colour, seed, sweep)
        transport.exchange()
        for s in shards:
            s.end(colour)
        return
    for s in shards:
        s.ctx.colour_pass(colour, seed, sweep)
    _serial_exchange(shards, transport, colour)


def _serial_exchange(shards: Sequence[Shard], transport, colour: int):
    """pack (main streams) -> records move (side streams) -> unpack (main streams), with host
    synchronisations in between: the round-1 protocol, kept for comparison and for set-up."""
    for s in shards:
        s.pack(colour)
    for s in shards:
        if s.side is not None:
            s.ctx.synchronize()
    transport.exchange()
    for s in shards:
        if s.side is not None:
            s.side.synchronize()
    for s in shards:
        s.unpack(colour)


def refresh_halos(shards: Sequence[Shard], transport):
    """Every shard's halo rows := its neighbours' boundary sites (after histories were created shard by
    shard, e.g. epv_simulate_adopt): the three boundary sites of either side travel as the records of
    colours 0, 1, 2.  The tri_llh cache is not valid afterwards (call reset)."""
    agree_halo_capacity(shards, transport, grow=False)
    for colour in range(3):
        _serial_exchange(shards, transport, colour)
    # boundary rows longer than half a record: grow before the first sweep
    agree_halo_capacity(shards, transport, grow=True)


def agree_halo_capacity(shards: Sequence[Shard], transport, grow: bool = True) -> int:
    """The halo-row capacity (hence the record size) must be the same on every shard: the maximum of
    what the shards have / ask for, set everywhere; records reallocated when it changed.  Synchronises
    every shard and (across ranks) the transport; returns the capacity."""
    want = 0
    for s in shards:
        cap, need = s.ctx.halo_capacity()
        want = max(want, cap, need if grow else 0)
    want = transport.allreduce_max(want) if hasattr(transport, "allreduce_max") else want
    for s in shards:
        s.ctx.halo_set_capacity(want)
        s.ensure_records()
    return want


class CapacityWatch:
    """The between-sweeps capacity check without stalling the GPUs.  agree_halo_capacity costs a host
    synchronisation, an all-reduce and a read-back per sweep: ~0.3 ms of idle GPU, x 20 sweeps = the whole
    weak-scaling loss measured at 8 GPUs (380.9 vs 373.8 ms per E-step) and 10 % of a strong-scaling step.
    Here the longest packed boundary row is copied on the device after a sweep's last pass
    (epv_halo_need_async), reduced over shards / ranks (max) on a stream of its own and copied to pinned
    memory; the host reads it at the end of the NEXT sweep -- long complete, no wait -- and only when twice
    that value passes the capacity does every rank (they all see the same reduced value) take the
    synchronising path and grow.  Rows are redrawn, not grown, by an update, so the one-sweep lag does not
    change what the half-full rule protects against; a row that outgrows a record still raises an error."""

    def __init__(self, shards: Sequence[Shard], transport):
        self.shards, self.transport = list(shards), transport
        self.pending, self.cap = False, 0
        self.ok = (all(hasattr(s.ctx, "halo_need_async") and s.side is not None for s in self.shards)
                   and hasattr(transport, "allreduce_max_tensor"))
        if self.ok:
            import torch
            self.need = [torch.zeros(1, dtype=torch.int64, device=s.device) for s in self.shards]
            self.stream = torch.cuda.Stream(device=self.shards[0].device)
            self.host = torch.zeros(1, dtype=torch.int64).pin_memory()
            self.event = torch.cuda.Event()

    def agree(self, grow: bool = True):
        self.cap = agree_halo_capacity(self.shards, self.transport, grow)

    def after_sweep(self):
        if not self.ok:
            return self.agree()
        import torch
        if self.pending:
            self.event.synchronize()                     # recorded a sweep ago
            if 2 * int(self.host[0]) > self.cap:
                self.agree()
        for s, t in zip(self.shards, self.need):
            s.side.wait_stream(self.stream)              # (the previous reduction read these words)
            s.ctx.halo_need_async(t.data_ptr(), s.side.cuda_stream)     # behind the sweep's last pack / unpack
            self.stream.wait_stream(s.side)
        with torch.cuda.stream(self.stream):
            total = self.need[0]
            for t in self.need[1:]:
                total = torch.maximum(total, t.to(total.device, non_blocking=True))
            total = self.transport.allreduce_max_tensor(total)
            self.host.copy_(total, non_blocking=True)
            self.event.record(self.stream)
        self.pending = True


def sweep(shards: Sequence[Shard], transport, seed: int, sweep_idx: int, overlap: bool = True, watch=None):
    """SingleSiteSampler::single_iteration over all shards (3 colour passes)."""
    for colour in range(3):
        colour_pass(shards, transport, colour, seed, sweep_idx, overlap)
    if watch is not None:
        watch.after_sweep()
    else:
        agree_halo_capacity(shards, transport)


def run_mcmc(shards: Sequence[Shard], transport, burn_in: int, batch: int, seed: int,
             first_sweep: int = 0, allreduce=None, overlap: bool = True):
    """SingleSiteSampler::run_mcmc (reference SingleSiteSampler.cpp:550-598) over shards.
    Returns batch-averaged J, D [n_nodes, 8] and the acceptance rate over all sites."""
    sw = first_sweep
    watch = CapacityWatch(shards, transport)
    watch.agree(grow=True)                                 # uploads size it per shard; growth a previous run saw late
    for _ in range(burn_in):
        sweep(shards, transport, seed, sw, overlap, watch)
        sw += 1
    for s in shards:
        s.ctx.accept_count(reset=True)
    n_nodes = shards[0].ctx.n_nodes
    n_global = shards[0].n_global
    k = n_nodes * 8
    if hasattr(shards[0].ctx, "stats_export"):
        # exact path: the batch's J/D passes add into integer accumulators on every shard; the shards'
        # accumulators (+ accept counts) are summed as int64 -- all-reduce across ranks -- and folded once,
        # so J, D and the accept rate do not depend on the number of shards, to the last bit
        import torch
        for s in shards:
            s.ctx.stats_begin()
        for _ in range(batch):
            sweep(shards, transport, seed, sw, overlap, watch)
            sw += 1
            for s in shards:
                s.ctx.stats_accumulate()
        words = shards[0].ctx.stats_words()
        total = None
        for s in shards:
            t = torch.zeros(words + 1, dtype=torch.int64, device=s.device)
            s.ctx.stats_export(t.data_ptr())
            t[words] = s.ctx.accept_count(reset=True)       # (synchronises the context's stream)
            total = t if total is None else total + t.to(total.device)
        if hasattr(transport, "allreduce_int"):
            total = transport.allreduce_int(total)
        torch.cuda.synchronize(total.device)
        J, D = shards[0].ctx.stats_fold(total.data_ptr())
        return J / batch, D / batch, float(total[words].item()) / (batch * (n_global - 2))
    Js, Ds = np.zeros((n_nodes, 8)), np.zeros((n_nodes, 8))
    for _ in range(batch):
        sweep(shards, transport, seed, sw, overlap, watch)
        sw += 1
        for s in shards:
            J, D = s.ctx.suffstats()
            Js += J
            Ds += D
    acc = float(sum(s.ctx.accept_count(reset=True) for s in shards))
    flat = np.concatenate([Js.ravel(), Ds.ravel(), [acc]])
    if allreduce is not None:
        flat = allreduce(flat)
    return (flat[:k].reshape(n_nodes, 8) / batch, flat[k:2 * k].reshape(n_nodes, 8) / batch,
            flat[-1] / (batch * (n_global - 2)))


def mcem_iteration(shards: Sequence[Shard], transport, burn_in: int, batch: int, seed: int,
                   first_sweep: int, rates, T, branches, optimize_branches: bool = False,
                   param_tol: float = 1e-10, allreduce=None, overlap: bool = True):
    """One MCEM iteration over shards -- the body of the loop at
    prog/epievo_est_params_histories.cpp:239-264, i.e. what epv_mcem_iteration does on one
    GPU: new model -> reset -> E-step (run_mcmc, J/D all-reduced) -> M-step on the host
    (identical on every rank) -> scale_jump_times.  Returns (rates, T, branches, baseline,
    acc_rate, llh)."""
    import ctypes as C

    from . import _lib as L
    lib = L.load()
    r = np.ascontiguousarray(rates, np.float64).copy()
    t = np.ascontiguousarray(T, np.float64).copy()
    br = np.ascontiguousarray(branches, np.float64).copy()
    bl = np.zeros(4)
    for s in shards:
        s.ctx.set_model(r, t)
        s.ctx.reset()
    J, D, acc = run_mcmc(shards, transport, burn_in, batch, seed, first_sweep, allreduce=allreduce, overlap=overlap)
    Jf, Df = np.ascontiguousarray(J.reshape(-1)), np.ascontiguousarray(D.reshape(-1))
    n_nodes = shards[0].ctx.n_nodes

    def d(a):
        return a.ctypes.data_as(L.dp)
    if not optimize_branches:
        llh = lib.epv_estimate_rates(n_nodes, d(Jf), d(Df), param_tol, d(r), d(t), d(bl))
        lib.epv_set_one_change(d(r), n_nodes, d(br))
    else:
        llh = lib.epv_estimate_rates_and_branches(n_nodes, d(Jf), d(Df), param_tol, d(r), d(br), d(t), d(bl))
    for s in shards:
        s.ctx.scale_jump_times(br)
    return r, t, br, bl, acc, llh


def gather_paths(parts: Sequence[PathSet], n_global: int, world: int) -> PathSet:
    """Owned blocks of every shard's local PathSet (halo stripped) -> one PathSet."""
    n_nodes = parts[0].n_nodes
    init = np.zeros((n_nodes, n_global), np.uint8)
    lens = np.zeros((n_nodes, n_global), np.int64)
    chunks = [[] for _ in range(n_nodes)]
    for r, ps in enumerate(parts):
        lo, hi, hl, hh = shard_bounds(n_global, world, r)
        own = ps.site_slice(hl, hl + (hi - lo))
        init[:, lo:hi] = own.init_state.reshape(n_nodes, -1)
        lens[:, lo:hi] = own.row_len()
        for b in range(n_nodes):
            a, e = int(own.row_off[b * own.n_sites]), int(own.row_off[(b + 1) * own.n_sites])
            chunks[b].append(own.jumps[a:e])
    off = np.zeros(n_nodes * n_global + 1, np.uint64)
    np.cumsum(lens.reshape(-1), out=off[1:].view(np.int64))
    jumps = np.concatenate([np.concatenate(c) for c in chunks]) if n_nodes else np.zeros(0)
    return PathSet(n_global, n_nodes, init.reshape(-1), off, jumps, parts[0].tot_time.copy(),
                   list(parts[0].node_names))
