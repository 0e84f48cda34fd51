"""N>1 host logic on CPU: two processes, torch.distributed gloo backend, the production
protocol code (epievo_b200.sharded: colour_pass / run_mcmc / DistExchange / all-reduce)
over CPU stand-in shards (tests/cpu_shard.py).  The 2-shard result must equal the
1-shard result exactly: paths, cached tri_llh, J/D and accept counts."""
import os
import socket
import sys

import numpy as np
import pytest
import torch.multiprocessing as mp

from epievo_b200 import sharded as S


def test_partition_is_colour_aligned_and_covers():
    for n, w in ((100, 3), (1001, 8), (12, 2), (10**8, 8)):
        blocks = S.partition(n, w)
        assert blocks[0][0] == 0 and blocks[-1][1] == n
        for (a, b), (c, d) in zip(blocks[:-1], blocks[1:]):
            assert b == c and c % 3 == 0
    with pytest.raises(ValueError):
        S.partition(20, 4)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, case, n, out_dir):
    here = os.path.dirname(os.path.abspath(__file__))
    for p in (os.path.dirname(here), here):
        if p not in sys.path:
            sys.path.insert(0, p)
    import torch.distributed as dist
    import cpu_libs
    from cpu_shard import CpuShard
    from epievo_b200.paths import Tree
    from epievo_b200.synthetic import simulate_paths
    from synth import TEST_PARAMS, ref_test_tree
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    tree = ref_test_tree() if case == "testtree" else Tree.pair(1.0)
    rates, T, _, _ = cpu_libs.oracle().model_from_params(*TEST_PARAMS)
    ps = simulate_paths(tree, n, rates, T, seed=9)
    sh = CpuShard(tree, rates, T, n, world, rank)
    sh.ctx.upload_paths(S.local_paths(ps, n, world, rank))
    tr = S.DistExchange(sh)
    J, D, acc = S.run_mcmc([sh], tr, 1, 2, seed=31, first_sweep=4, allreduce=tr.allreduce_sum)
    parts = [None]*world
    dist.all_gather_object(parts, (sh.ctx.download_paths(), sh.ctx.tri_llh()))
    if rank == 0:
        got = S.gather_paths([p for p, _ in parts], n, world)
        one = CpuShard(tree, rates, T, n, 1, 0)
        one.ctx.upload_paths(ps)

        class NoExchange:
            def exchange(self):
                pass
        J1, D1, acc1 = S.run_mcmc([one], NoExchange(), 1, 2, seed=31, first_sweep=4)
        want, tri1 = one.ctx.download_paths(), one.ctx.tri_llh()
        ok = got.same_as(want) and np.array_equal(J, J1) and np.allclose(D, D1, rtol=1e-12) and acc == acc1
        for r, (p, tri) in enumerate(parts):
            lo, hi, hl, hh = S.shard_bounds(n, world, r)
            ok = ok and np.array_equal(tri[hl + 1:hl + (hi - lo) - 1], tri1[lo + 1:hi - 1])
            ok = ok and p.same_as(S.local_paths(want, n, world, r))       # halos are current
        open(os.path.join(out_dir, "result"), "w").write("ok" if ok else "mismatch")
    dist.destroy_process_group()


@pytest.mark.parametrize("case,n", [("testtree", 120), ("pair", 90)])
def test_two_gloo_ranks_equal_one(tmp_path, case, n):
    port = _free_port()
    mp.spawn(_worker, args=(2, port, case, n, str(tmp_path)), nprocs=2, join=True)
    assert open(tmp_path / "result").read() == "ok"
