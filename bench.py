#!/usr/bin/env python
"""bench.py -- MH site x branch updates per second of the MCMC E-step.

A "step" is one call of the reference's SingleSiteSampler::run_mcmc with its default
schedule (burn-in L=10 sweeps, then batch B=10 x (sweep + per-branch J/D pass);
reference prog/epievo_est_params_histories.cpp:104-106, SingleSiteSampler.cpp:550-598)
over one GPU's block of sites.  Updates per step = (L+B) * (n_sites-2) * n_branches.

  value : paths already resident in HBM when the timed region starts (epv_run_mcmc)
  e2e   : the same call with HOST buffers -- pinned host paths uploaded, E-step run,
          paths + J/D read back, all inside the timed region
  roofline : the colour-pass update kernel (mh_update_kernel), algorithmic bytes per launch
          = (4/3) * n_sites * n_branches * (4 + 1/8 + 8*jbar)   
          over its mean launch time from CUDA events around every launch
  cpu_baseline : the unmodified reference (oracle/_ref) on one host core, bounded sample

python bench.py --impl reference times the reference's CPU path on all host cores.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "tests")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

METRIC = "mh_site_branch_updates_per_sec"
UNIT = "updates/s"
DEFAULT_SITES = {"bal8": 12_500_000, "testtree": 30_000_000, "pairwise": 1_000_000,
                 "bal32": 3_000_000}
WORKLOAD_DESC = {
    "bal8": "north-star headline shard: balanced 8-leaf tree (14 branches x 0.05), "
            "10^8 sites / 8 GPUs = 1.25e7 sites per GPU",
    "testtree": "BASELINE config 4: reference test tree (4 branches), 3e7 sites",
    "pairwise": "BASELINE config 2: single branch T=1.0, 1e6 sites",
    "bal32": "BASELINE config 5 shard: balanced 32-leaf tree (62 branches x 0.05)",
}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms during the timed region."""
    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index = index
        self.proc = None
        self.path = None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.FIELDS,
                 "--format=csv,noheader,nounits", "-lms", "200"],
                stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None
        self._start_nvml()

    # A timed region shorter than nvidia-smi's first report (strong scaling at N=8: 3 steps of 57 ms) would end
    # with no sample at all: the same counters are also read in-process through NVML every 10 ms.
    def _start_nvml(self):
        self.nvml, self.nvml_stop, self.nvml_thread = [], threading.Event(), None
        try:
            import pynvml
            pynvml.nvmlInit()
            h = pynvml.nvmlDeviceGetHandleByIndex(self.index)
            mx = float(pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM))
            bits = (("hw_slowdown", pynvml.nvmlClocksEventReasonHwSlowdown),
                    ("hw_thermal_slowdown", pynvml.nvmlClocksEventReasonHwThermalSlowdown),
                    ("sw_thermal_slowdown", pynvml.nvmlClocksEventReasonSwThermalSlowdown),
                    ("sw_power_cap", pynvml.nvmlClocksEventReasonSwPowerCap))

            def loop():
                while not self.nvml_stop.is_set():
                    try:
                        sm = float(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM))
                        r = int(pynvml.nvmlDeviceGetCurrentClocksEventReasons(h))
                        self.nvml.append((sm, mx, [nm for nm, b in bits if r & b]))
                    except Exception:
                        pass
                    self.nvml_stop.wait(0.01)

            self.nvml_thread = threading.Thread(target=loop, daemon=True)
            self.nvml_thread.start()
        except Exception:
            self.nvml_thread = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if getattr(self, "nvml_thread", None) is not None:
            self.nvml_stop.set()
            self.nvml_thread.join(timeout=2)
        if self.proc is not None:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except Exception:
                self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        try:
            for line in open(self.path):
                tok = [t.strip() for t in line.split(",")]
                if len(tok) < 7:
                    continue
                try:
                    sm.append(float(tok[0]))
                    mx.append(float(tok[1]))
                except ValueError:
                    continue
                for nm, val in zip(names, tok[3:7]):
                    if val.lower().startswith("active"):
                        reasons.add(nm)
            os.unlink(self.path)
        except Exception:
            pass
        out["source"] = "nvidia-smi"
        if not sm and self.nvml:
            out["source"] = "nvml"
            for a, b, r in self.nvml:
                sm.append(a); mx.append(b); reasons.update(r)
        if sm:
            sm.sort()
            out.update(sm_mhz=sm[len(sm) // 2], sm_max_mhz=max(mx), samples=len(sm))
        out["reasons"] = sorted(reasons)
        return out


def make_tree(workload):
    """The workload's tree from the product's own containers (no oracle/ on the b200 arm)."""
    from epievo_b200.paths import Tree, parse_newick
    if workload == "bal8":
        return Tree.balanced(8, 0.05)
    if workload == "bal32":
        return Tree.balanced(32, 0.05)
    if workload == "pairwise":
        return Tree.pair(1.0)
    return parse_newick("((C:0.03,D:0.06)E:0.02,F:0.1)G:0.0;")     # reference test/tree.nwk


def test_model():
    """reference test/test.param through the library's epv_model_from_params (rates scaled)."""
    import ctypes as C

    import numpy as np

    from epievo_b200 import _lib as L
    rates, T, Q, bl = np.zeros(8), np.zeros(4), np.zeros(4), np.zeros(4)
    d = lambda a: a.ctypes.data_as(L.dp)
    L.load().epv_model_from_params(0.844912, 0.893359, -0.8, -1.8, 1, d(rates), d(T), d(Q), d(bl))
    return rates, T


def run_reference(args):
    """--impl reference: the unmodified reference CPU path on all host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import cpu_bench
    procs = args.cpu_procs or (os.cpu_count() or 1)
    n_s = args.cpu_sites
    tree = cpu_bench.make_tree(args.workload)
    n_br = tree.n_nodes - 1
    sweeps = args.burn_in + args.batch
    kind, per_step = cpu_bench.run(args.workload, n_s, args.burn_in, args.batch,
                                   args.warmup + args.steps, procs)
    timed = per_step[args.warmup:]
    total_s = sum(timed)
    updates = procs * (n_s - 2) * n_br * sweeps * args.steps
    value = updates / total_s
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT,
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * total_s / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(args, n_br, sample="%d processes x %d sites" % (procs, n_s)),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": procs, "kind": kind,
                         "sample": "%d independent processes x %d-site blocks of the same workload, "
                                   "each step one run_mcmc(L=%d,B=%d)" % (procs, n_s, args.burn_in, args.batch)},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))
    return 0


def workload_config(args, n_br, sample=None, n_global=None, world=1, gen="host"):
    desc = WORKLOAD_DESC[args.workload]
    if args.sites != DEFAULT_SITES[args.workload] or args.scaling == "strong":
        desc += " [run at --sites %d, %s scaling]" % (args.sites, args.scaling)
    per_gpu = args.sites if args.scaling == "weak" else (n_global or args.sites) // max(1, world)
    cfg = {"workload": args.workload, "description": desc,
           "n_sites_per_gpu": per_gpu, "n_sites_total": n_global or args.sites, "n_branches": n_br,
           "histories": "numpy generator + upload" if gen == "host" else "epv_simulate on the device, adopted in place",
           "burn_in": args.burn_in,
           "batch": args.batch, "step": "one run_mcmc E-step = (L+B) sweeps + B J/D passes",
           "parallelism": "contiguous site blocks, one per GPU; 3-site halos, one NCCL send/recv pair "
                          "per neighbour per colour pass (edge groups first on a side stream, the exchange "
                          "overlaps the interior update), one int64 all-reduce of the J/D accumulators per E-step",
           "l2": "inputs (>=0.5 GB of rows per pass) are larger than the 126 MB L2"}
    if sample:
        cfg["sample"] = sample
    return cfg


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="bal8", choices=sorted(DEFAULT_SITES))
    ap.add_argument("--sites", type=int, default=0, help="sites per GPU (default: per workload)")
    ap.add_argument("--burn-in", type=int, default=10)
    ap.add_argument("--batch", type=int, default=10)
    ap.add_argument("--cpu-sites", type=int, default=25_000,
                    help="sites per process of the CPU sample")
    ap.add_argument("--cpu-procs", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-mcem", action="store_true")
    ap.add_argument("--scaling", choices=("weak", "strong"), default="weak",
                    help="weak: --sites per GPU; strong: --sites in total, split over the GPUs")
    ap.add_argument("--gen", choices=("auto", "host", "device"), default="auto",
                    help="synthetic histories from the numpy generator (host) or epv_simulate on the GPU "
                         "(device: shards too large to stage through host arrays; no e2e leg); "
                         "auto = device above 4e8 rows per GPU")
    ap.add_argument("--optimize-branches", action="store_true",
                    help="the MCEM iteration updates branch lengths too (the reference's -b)")
    ap.add_argument("--serial-exchange", action="store_true",
                    help="N>1: the round-1 protocol (pass, pack, exchange, unpack on one stream) "
                         "instead of the overlapped one")
    ap.add_argument("--exchange", choices=("compact", "csr"), default="compact",
                    help="host exchange format of the e2e leg (u16 row codes, or 9-byte CSR rows)")
    args = ap.parse_args()
    if not args.sites:
        args.sites = DEFAULT_SITES[args.workload]
    if args.impl == "reference":
        return run_reference(args)

    import numpy as np
    import torch
    import torch.distributed as dist

    from epievo_b200.device import Context
    from epievo_b200.synthetic import simulate_paths

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the B200 path has no CPU fallback "
                         "(use --impl reference for the CPU reference arm)")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    tree = make_tree(args.workload)
    rates, T = test_model()
    n_br = tree.n_nodes - 1
    sweeps = args.burn_in + args.batch
    from epievo_b200 import sharded as S
    from epievo_b200.paths import PathSet
    n_global = args.sites * world if args.scaling == "weak" else args.sites
    lo, hi, halo_lo, halo_hi = S.shard_bounds(n_global, world, rank)
    n_own = hi - lo
    n = n_own + halo_lo + halo_hi           # local sites: owned block + halos
    n_rows = n * tree.n_nodes
    gen = args.gen if args.gen != "auto" else ("device" if n_rows > 4e8 else "host")

    ctx = Context(local)
    stream = torch.cuda.Stream()           # a real (non-default) stream the context launches on
    torch.cuda.set_stream(stream)
    ctx.set_stream(stream.cuda_stream)
    ctx.set_tree(tree)
    ctx.set_model(rates, T)
    shard = transport = None
    if world > 1:
        shard = S.Shard(ctx, n_global, world, rank, torch.device("cuda", local))
        transport = S.DistExchange(shard)
    tot_time = np.array(tree.branches, np.float64)
    tot_time[0] = 0.0
    h_init = h_off = h_jmp = h_code = None
    jcap = nj_host = 0
    if gen == "host":
        ps = simulate_paths(tree, n_own, rates, T, seed=12345 + rank)
        if world > 1:
            # neighbours' boundary sites become this rank's halos (gathered once, at setup)
            edges = [None] * world
            dist.all_gather_object(edges, (ps.site_slice(0, S.HALO), ps.site_slice(n_own - S.HALO, n_own)))
            parts = ([edges[rank - 1][1]] if halo_lo else []) + [ps] + ([edges[rank + 1][0]] if halo_hi else [])
            ps = PathSet.concat_sites(parts)
        assert ps.n_sites == n
        # pinned host buffers in the C-ABI exchange format (jumps sized with headroom)
        jcap = int(ps.jumps.size * 1.5) + 65536
        h_init = torch.empty(n_rows, dtype=torch.uint8, pin_memory=True)
        h_off = torch.empty(n_rows + 1, dtype=torch.int64, pin_memory=True)
        h_jmp = torch.empty(jcap, dtype=torch.float64, pin_memory=True)
        h_init.numpy()[:] = ps.init_state
        h_off.numpy()[:] = ps.row_off.view(np.int64)
        h_jmp.numpy()[:ps.jumps.size] = ps.jumps
        h_code = torch.empty(n_rows, dtype=torch.int16, pin_memory=True)   # compact format: u16 codes
        nj_host = int(ps.jumps.size)
        tot_time = ps.tot_time.copy()
        del ps
        ctx.upload_raw(n, tree.n_nodes, h_init.data_ptr(), h_off.data_ptr(), h_jmp.data_ptr(), tot_time)
    else:
        # shards too large to stage through host arrays (config 5: 7.9e8 rows per GPU): the library's own
        # forward simulator writes the histories on the device and the context adopts them; every rank
        # simulates its local sites and then takes its halo rows from the neighbours
        ctx.simulate_adopt(n, 12345 + rank)
        if world > 1:
            S.refresh_halos([shard], transport)
    ctx.reset()

    def e_step(first_sweep):
        """One E-step = run_mcmc(L, B).  N=1: the C-ABI call; N>1: the same schedule with a
        halo exchange after every colour pass and one all-reduce of the J/D sums."""
        if world == 1:
            return ctx.run_mcmc(args.burn_in, args.batch, seed=2024, first_sweep=first_sweep)
        return S.run_mcmc([shard], transport, args.burn_in, args.batch, seed=2024,
                          first_sweep=first_sweep, allreduce=transport.allreduce_sum,
                          overlap=not args.serial_exchange)

    sweep_no = 0
    for _ in range(args.warmup):
        e_step(sweep_no)
        sweep_no += sweeps

    # ---- timed region 1: device-resident E-steps ------------------------------------
    ctx.enable_kernel_timing(True)
    ctx.mh_kernel_time(reset=True)
    clocks = ClockSampler(local)
    launches0 = ctx.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    clocks.start()
    e0.record(stream)
    acc = 0.0
    for _ in range(args.steps):
        _, _, acc = e_step(sweep_no)
        sweep_no += sweeps
    e1.record(stream)
    barrier()
    clk = clocks.stop()
    ms = e0.elapsed_time(e1)
    mh_ms, mh_launches = ctx.mh_kernel_time(reset=True)
    rebuild_ms = ctx.rebuild_kernel_time()
    ctx.enable_kernel_timing(False)
    launches = ctx.launch_count() - launches0
    jc = ctx.jump_counts()
    jbar = float(jc.sum()) / (n_own * n_br)

    t = torch.tensor([ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_max = float(t.item())
    updates_per_step = (n_global - 2) * n_br * sweeps
    value = updates_per_step * args.steps / (ms_max * 1e-3)

    # ---- timed region 2: end to end through the C-ABI with host buffers --------------
    e2e = None
    if not args.no_e2e and gen == "host":
        n_sites_, n_nodes_, nj = ctx.paths_size()
        if nj > jcap:
            raise SystemExit("bench.py: pinned jump buffer too small")
        compact = args.exchange == "compact"
        if compact:
            nj_host = ctx.download_compact_raw(h_code.data_ptr(), h_jmp.data_ptr(), jcap)
        else:
            ctx.download_raw(h_init.data_ptr(), h_off.data_ptr(), h_jmp.data_ptr())
        h2d = d2h = 0
        step_ev = []
        # one untimed pass of the same three calls first (the first compact upload sizes buffers and loads its
        # kernels), then args.steps timed ones, an event after each
        for it in range(1 + args.steps):
            if it == 1:
                h2d = d2h = 0
                barrier()
                e0.record(stream)
            if compact:
                nj_in = nj_host
                ctx.upload_compact_raw(n, tree.n_nodes, h_code.data_ptr(), h_jmp.data_ptr(), nj_in, tot_time)
            else:
                nj_in = int(h_off[n_rows].item())
                ctx.upload_raw(n, tree.n_nodes, h_init.data_ptr(), h_off.data_ptr(), h_jmp.data_ptr(), tot_time)
            J, D, acc2 = e_step(sweep_no)
            sweep_no += sweeps
            if compact:
                nj_host = nj_out = ctx.download_compact_raw(h_code.data_ptr(), h_jmp.data_ptr(), jcap)
            else:
                _, _, nj_out = ctx.paths_size()
                if nj_out > jcap:
                    raise SystemExit("bench.py: pinned jump buffer too small")
                ctx.download_raw(h_init.data_ptr(), h_off.data_ptr(), h_jmp.data_ptr())
            per_row = 2 * n_rows if compact else n_rows + 8 * (n_rows + 1)
            h2d += per_row + 8 * nj_in
            d2h += per_row + 8 * nj_out + 2 * 8 * 8 * tree.n_nodes
            if it >= 1:
                step_ev.append(torch.cuda.Event(enable_timing=True))
                step_ev[-1].record(stream)
        e1.record(stream)
        barrier()
        t2 = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t2, op=dist.ReduceOp.MAX)
        e2e = {"value": updates_per_step * args.steps / (float(t2.item()) * 1e-3), "unit": UNIT,
               "h2d_bytes_per_step": h2d // args.steps, "d2h_bytes_per_step": d2h // args.steps,
               "ms_per_step": float(t2.item()) / args.steps,
               "ms_each_step": [round(a.elapsed_time(b), 3) for a, b in zip([e0] + step_ev[:-1], step_ev)],
               "untimed_warmup_steps": 1,
               "call": ("epv_upload_paths_compact + epv_run_mcmc + epv_download_paths_compact"
                        if compact else "epv_upload_paths + epv_run_mcmc + epv_download_paths")
                       + " (pinned host buffers)"}

    # ---- roofline of the dominant kernel -----------------------------------------------
    peaks_file = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_file):
        peak, peak_src = float(json.load(open(peaks_file))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    else:
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    row_bytes = 4.0 + 0.125 + 8.0 * jbar
    bytes_per_launch = (4.0 / 3.0) * n_own * n_br * row_bytes
    avg_launch_ms = mh_ms / max(1, mh_launches)
    achieved = bytes_per_launch / (avg_launch_ms * 1e-3) / 1e9 if avg_launch_ms > 0 else 0.0
    traffic = None                      # dram bytes of the kernel from the committed ncu capture
    traffic_source = issue_pct = lanes = None
    for tname in ("r02_traffic.json", "r01_traffic.json"):      # the newest committed capture
        tfile = os.path.join(ROOT, "profiles", tname)
        if not os.path.exists(tfile):
            continue
        tj = json.load(open(tfile))
        if tj.get("workload") == args.workload and tj.get("n_sites_per_gpu") == args.sites:
            traffic = tj["traffic_bytes_per_launch"]
            traffic_source = ("ncu --set full capture of the same kernel and workload, committed as profiles/%s; "
                              "NOT measured in this run" % tname)
            issue_pct, lanes = tj.get("issue_active_pct"), tj.get("lanes_per_inst")
        break
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_source,
                # what actually bounds the kernel (same capture): issue slots used, active lanes per instruction
                "issue_active_pct": issue_pct, "lanes_per_inst": lanes,
                "kernel": "mh_update_kernel",
                "peak_source": peak_src, "avg_launch_ms": avg_launch_ms, "launches_timed": mh_launches,
                "algorithmic_bytes_per_launch": bytes_per_launch, "jumps_per_row": jbar,
                "kernel_share_of_step": mh_ms / ms if ms > 0 else None,
                "rebuild_kernel_avg_ms": rebuild_ms / max(1, mh_launches),
                "rebuild_kernel_share_of_step": rebuild_ms / ms if ms > 0 else None}

    # ---- the J/D pass on its own (epv_suffstats: kernel + 1 KB read-back), HBM roofline ----
    sst0, sst1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ctx.suffstats()
    sst0.record(stream)
    for _ in range(5):
        ctx.suffstats()
    sst1.record(stream)
    torch.cuda.synchronize()
    ss_ms = sst0.elapsed_time(sst1) / 5
    ss_bytes = n_own * n_br * row_bytes
    suffstats = {"ms": ss_ms, "algorithmic_bytes": ss_bytes, "achieved_gbs": ss_bytes / (ss_ms * 1e-3) / 1e9,
                 "frac_of_hbm_peak": ss_bytes / (ss_ms * 1e-3) / 1e9 / peak,
                 "rows_per_s": n_own * n_br / (ss_ms * 1e-3)}

    # ---- one whole MCEM iteration, paths resident (reset + E-step + M-step + rescale) ------
    # prog/epievo_est_params_histories.cpp:239-264; N=1: epv_mcem_iteration, N>1: the same
    # steps with the halo exchange and one all-reduce (sharded.mcem_iteration).  Wall clock
    # between device synchronisations (the M-step runs on the host), max over ranks.
    def mcem_once(first_sweep):
        if world == 1:
            return ctx.mcem_iteration(args.burn_in, args.batch, 2024, first_sweep, rates, T, tree.branches,
                                      optimize_branches=args.optimize_branches)
        return S.mcem_iteration([shard], transport, args.burn_in, args.batch, 2024, first_sweep, rates, T,
                                tree.branches, optimize_branches=args.optimize_branches,
                                allreduce=transport.allreduce_sum, overlap=not args.serial_exchange)
    mcem = None
    if not args.no_mcem:
        barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        out = mcem_once(sweep_no)
        torch.cuda.synchronize()
        dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        sweep_no += sweeps
        mcem = {"ms": float(dt.item()) * 1e3, "burn_in": args.burn_in, "batch": args.batch,
                "sites": int(n_global), "acc_rate": float(out[4]), "optimize_branches": bool(args.optimize_branches),
                "peak_device_memory_gb": (torch.cuda.mem_get_info()[1] - torch.cuda.mem_get_info()[0]) / 1e9,
                "what": "set_model + reset + run_mcmc(L,B) + M-step (estimate_rates, host) + scale_jump_times"}

    # ---- CPU baseline (rank 0, N=1 only) ---------------------------------------------------
    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        n_s = 50_000
        bi, ba = 5, 5
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import cpu_bench                  # the CPU leg only: the checker timed, never part of the GPU arm
        kind, per_step = cpu_bench.run(args.workload, n_s, bi, ba, 1, 1)
        cpu_baseline = {"value": (n_s - 2) * n_br * (bi + ba) / per_step[0], "unit": UNIT, "cores": 1,
                        "kind": kind,
                        "sample": "%d sites of the same workload, one run_mcmc(L=%d,B=%d), single "
                                  "thread (the reference is single-threaded), %.1f s" % (n_s, bi, ba, per_step[0])}

    if mcem and cpu_baseline:
        mcem["cpu_reference_est_s"] = (n_global - 2) * n_br * sweeps / cpu_baseline["value"]
        mcem["cpu_note"] = "reference on 1 core, extrapolated from the cpu_baseline sample (its cost is linear in sites)"
    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_max / args.steps, "higher_is_better": True,
            "scaling": args.scaling, "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(args, n_br, n_global=n_global, world=world, gen=gen), "clocks": clk, "e2e": e2e,
            "gpu_launches": int(launches), "roofline": roofline, "cpu_baseline": cpu_baseline,
            "acc_rate": acc, "suffstats_pass": suffstats, "mcem_iter": mcem,
        }
        print(json.dumps(line))
    ctx.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


def _quiet_stdout():
    """stdout must carry exactly ONE line, the JSON record: libraries write there at will (NCCL prints its
    version on rank 0 at the first collective).  File descriptor 1 is pointed at stderr for the whole run and
    Python's sys.stdout at a copy of the original descriptor, so only this script's own print reaches it."""
    sys.stdout.flush()
    keep = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(keep, "w", buffering=1)


if __name__ == "__main__":
    _quiet_stdout()
    rc = main()
    sys.stdout.flush()
    sys.exit(rc)
