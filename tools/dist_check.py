"""torchrun check: N ranks (one GPU each, NCCL halo exchange) reproduce the single-GPU
run bit for bit.  Rank 0 also runs the unsharded problem and compares."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "oracle")):
    sys.path.insert(0, p)

import cpu_bench                                   # noqa: E402
from epievo_b200 import sharded as S               # noqa: E402
from epievo_b200.device import Context             # noqa: E402
from epievo_b200.synthetic import simulate_paths   # noqa: E402


def main():
    rank, world, local = (int(os.environ[k]) for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ok = True
    for wl, n in (("testtree", 30011), ("bal8", 9001), ("pairwise", 4000)):
        tree = cpu_bench.make_tree(wl)
        rates, T = cpu_bench.test_model()
        ps = simulate_paths(tree, n, rates, T, seed=5)          # same on every rank
        st = torch.cuda.Stream()
        torch.cuda.set_stream(st)
        ctx = Context(local)
        ctx.set_stream(st.cuda_stream)
        ctx.set_tree(tree); ctx.set_model(rates, T)
        sh = S.Shard(ctx, n, world, rank, torch.device("cuda", local))
        ctx.upload_paths(S.local_paths(ps, n, world, rank)); ctx.reset()
        tr = S.DistExchange(sh)
        J, D, acc = S.run_mcmc([sh], tr, 3, 4, seed=11, first_sweep=2, allreduce=tr.allreduce_sum)
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
            same = got.same_as(want) and np.array_equal(J, J1) and np.allclose(D, D1, rtol=1e-12) and acc == acc1
            print("%s n=%d world=%d: %s (acc %.4f)" % (wl, n, world, "BIT-EXACT" if same else "MISMATCH", acc))
            ok = ok and same
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
