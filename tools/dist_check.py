"""torchrun check: N ranks (one GPU each, NCCL halo exchange) reproduce the single-GPU run bit for bit --
with the overlapped protocol (edge groups first, records on the side stream) and with the serial one, on
the plain E-step and on the -b MCEM loop (estimate_rates_and_branches + scale_jump_times every iteration,
prog/epievo_est_params_histories.cpp:259-264).  Rank 0 also runs the unsharded problem and compares:
histories, J (exact), D (exact: integer accumulators all-reduced as int64), accept rate, and for -b the
parameters and branch lengths.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/dist_check.py
"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)

import bench                                       # noqa: E402  (make_tree / test_model of the b200 arm)
from epievo_b200 import sharded as S               # noqa: E402
from epievo_b200.device import Context             # noqa: E402
from epievo_b200.synthetic import simulate_paths   # noqa: E402


def make_shard(tree, rates, T, ps, n, world, rank, local):
    st = torch.cuda.Stream()
    torch.cuda.set_stream(st)
    ctx = Context(local)
    ctx.set_stream(st.cuda_stream)
    ctx.set_tree(tree); ctx.set_model(rates, T)
    sh = S.Shard(ctx, n, world, rank, torch.device("cuda", local))
    ctx.upload_paths(S.local_paths(ps, n, world, rank)); ctx.reset()
    return ctx, sh, S.DistExchange(sh), st


def main():
    rank, world, local = (int(os.environ[k]) for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ok = True
    rates, T = bench.test_model()
    for wl, n in (("testtree", 30011), ("bal8", 9001), ("pairwise", 4000), ("bal32", 2003)):
        tree = bench.make_tree(wl)
        ps = simulate_paths(tree, n, rates, T, seed=5)          # same on every rank
        for overlap in (True, False):
            ctx, sh, tr, st = make_shard(tree, rates, T, ps, n, world, rank, local)
            J, D, acc = S.run_mcmc([sh], tr, 3, 4, seed=11, first_sweep=2, allreduce=tr.allreduce_sum, overlap=overlap)
            parts = [None] * world
            dist.all_gather_object(parts, ctx.download_paths())
            ctx.close()
            if rank == 0:
                got = S.gather_paths(parts, n, world)
                one = Context(local)
                one.set_tree(tree); one.set_model(rates, T); one.upload_paths(ps); one.reset()
                J1, D1, acc1 = one.run_mcmc(3, 4, seed=11, first_sweep=2)
                want = one.download_paths()
                one.close()
                same = got.same_as(want) and np.array_equal(J, J1) and np.array_equal(D, D1) and acc == acc1
                print("%s n=%d world=%d %s: %s (acc %.4f)" % (wl, n, world, "overlapped" if overlap else "serial",
                                                             "BIT-EXACT" if same else "MISMATCH", acc))
                ok = ok and same
        # ---- the -b loop: 3 MCEM iterations with branch-length updates -----------------------------------
        if wl in ("testtree", "bal32"):
            ctx, sh, tr, st = make_shard(tree, rates, T, ps, n, world, rank, local)
            r, t, br = np.array(rates), np.array(T), np.array(tree.branches, np.float64)
            sweep = 0
            for _ in range(3):
                r, t, br, bl, acc, llh = S.mcem_iteration([sh], tr, 3, 3, 77, sweep, r, t, br, optimize_branches=True,
                                                          allreduce=tr.allreduce_sum)
                sweep += 6
            parts = [None] * world
            dist.all_gather_object(parts, ctx.download_paths())
            ctx.close()
            if rank == 0:
                got = S.gather_paths(parts, n, world)
                one = Context(local)
                one.set_tree(tree); one.set_model(rates, T); one.upload_paths(ps)
                r1, t1, br1 = np.array(rates), np.array(T), np.array(tree.branches, np.float64)
                sweep = 0
                for _ in range(3):
                    r1, t1, br1, bl1, acc1, llh1 = one.mcem_iteration(3, 3, 77, sweep, r1, t1, br1, optimize_branches=True)
                    sweep += 6
                want = one.download_paths()
                one.close()
                same = (got.same_as(want) and np.array_equal(r, r1) and np.array_equal(br, br1) and acc == acc1
                        and llh == llh1)
                print("%s n=%d world=%d -b MCEM x3: %s (branch[1] %.6g, llh %.6f)" % (
                    wl, n, world, "BIT-EXACT" if same else "MISMATCH", br[1], llh))
                ok = ok and same
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.broadcast(flag, 0)
    dist.destroy_process_group()
    sys.exit(0 if int(flag.item()) else 1)


if __name__ == "__main__":
    main()
