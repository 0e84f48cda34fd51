"""High-powered chain-vs-chain probe of the sampler's stationary statistics: R independent chains
of the reference's sequential sampler (oracle/_ref, mt19937) against R chains of the GPU's
colour-class sampler from the same starting paths; Welch z per (branch, triplet) batch-mean J and D
entry, mean z^2 over the entries (1.0 under the null when the entries are independent).

    python tools/stats_probe.py --arm cpu --R 512 --burn 30     # writes tools/_cache/stats_probe_cpu_*.npz
    python tools/stats_probe.py --arm gpu --R 512 --burn 30     # reads it, runs the GPU arm, prints the table

The CPU arm is deterministic (seeds 1000+c) and needs no GPU, so it is computed once (on all host
cores) and cached; tools/_cache/ is git-ignored but travels to the GPU box.
"""
import argparse
import os
import sys
from multiprocessing import Pool

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np

from epievo_b200.paths import Tree
from epievo_b200.synthetic import simulate_paths

TEST_PARAMS = (0.844912, 0.893359, -0.8, -1.8)
BATCH = 20


def problem(case, n):
    import cpu_libs
    rates, T, _, _ = cpu_libs.oracle().model_from_params(*TEST_PARAMS)
    tree = {"bal8": lambda: Tree.balanced(8, 0.05), "pair": lambda: Tree.pair(1.0)}[case]()
    return tree, rates, T, simulate_paths(tree, n, rates, T, seed=77)


def cpu_chunk(args):
    case, n, burn, chains = args
    import cpu_libs
    cpu = cpu_libs.best_cpu()
    tree, rates, T, ps = problem(case, n)
    out = []
    for c in chains:
        cpu.load_paths(ps); cpu.seed(1000 + c)
        J, D, _, _ = cpu.run_mcmc(tree, rates, T, burn, BATCH)
        out.append(np.concatenate([J[1:].ravel(), D[1:].ravel()]))
    return np.array(out)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--arm", choices=["cpu", "gpu"], required=True)
    ap.add_argument("--case", default="bal8")
    ap.add_argument("--n", type=int, default=600)
    ap.add_argument("--R", type=int, default=512)
    ap.add_argument("--burn", type=int, default=30)
    ap.add_argument("--procs", type=int, default=os.cpu_count())
    a = ap.parse_args()
    cache = os.path.join(ROOT, "tools", "_cache")
    os.makedirs(cache, exist_ok=True)
    f = os.path.join(cache, "stats_probe_cpu_%s_n%d_R%d_burn%d.npz" % (a.case, a.n, a.R, a.burn))
    if a.arm == "cpu":
        chunks = [list(range(k, a.R, a.procs)) for k in range(a.procs)]
        with Pool(a.procs) as pool:
            parts = pool.map(cpu_chunk, [(a.case, a.n, a.burn, ch) for ch in chunks])
        ref = np.zeros((a.R, parts[0].shape[1]))
        for ch, p in zip(chunks, parts):
            ref[ch] = p
        np.savez_compressed(f, ref=ref)
        print("wrote", f, ref.shape)
        return
    ref = np.load(f)["ref"]
    from epievo_b200.device import Context
    tree, rates, T, ps = problem(a.case, a.n)
    ctx = Context(0); ctx.set_tree(tree); ctx.set_model(rates, T)
    gpu = []
    for c in range(a.R):
        ctx.upload_paths(ps); ctx.reset()
        J, D, _ = ctx.run_mcmc(a.burn, BATCH, seed=5000 + c)
        gpu.append(np.concatenate([J[1:].ravel(), D[1:].ravel()]))
    gpu = np.array(gpu)
    R = a.R
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    np.savez_compressed(os.path.join(ROOT, "gpurun_out", "stats_probe_gpu_%s_n%d_R%d_burn%d.npz" % (a.case, a.n, R, a.burn)), gpu=gpu)
    z = (ref.mean(0) - gpu.mean(0))/np.sqrt(ref.var(0, ddof=1)/R + gpu.var(0, ddof=1)/R + 1e-300)
    print("case %s n=%d R=%d burn-in %d batch %d" % (a.case, a.n, R, a.burn, BATCH))
    for i in np.argsort(-np.abs(z))[:12]:
        print(i, "z=%.2f" % z[i], "ref=%.5g gpu=%.5g" % (ref[:, i].mean(), gpu[:, i].mean()))
    # null calibration of mean z^2 with the entries' actual correlation: split the CPU arm in two halves
    h = R//2
    zz = (ref[:h].mean(0) - ref[h:2*h].mean(0))/np.sqrt(ref[:h].var(0, ddof=1)/h + ref[h:2*h].var(0, ddof=1)/h + 1e-300)
    # effective number of independent entries from the correlation matrix of the pooled residuals
    X = np.concatenate([ref - ref.mean(0), gpu - gpu.mean(0)])
    X = X[:, X.std(0) > 0]
    C = np.corrcoef(X.T)
    k_eff = C.shape[0]**2/np.sum(C**2)
    print("mean z^2 = %.4f over %d entries (max |z| %.2f); effective independent entries %.1f => null sd of mean z^2 = %.3f"
          % (np.nanmean(z**2), z.size, np.abs(z).max(), k_eff, np.sqrt(2.0/k_eff)))
    # aggregated statistics (far more power than any single entry): jumps per sweep, overall and per branch
    k = ref.shape[1]//2
    def welch(x, y):
        return (x.mean() - y.mean())/np.sqrt(x.var(ddof=1)/len(x) + y.var(ddof=1)/len(y))
    print("total J per sweep: ref %.4f gpu %.4f  z=%.2f" % (ref[:, :k].sum(1).mean(), gpu[:, :k].sum(1).mean(),
                                                             welch(ref[:, :k].sum(1), gpu[:, :k].sum(1))))
    zb = [welch(ref[:, 8*b:8*b + 8].sum(1), gpu[:, 8*b:8*b + 8].sum(1)) for b in range(k//8)]
    print("per-branch total J, Welch z:", " ".join("%.2f" % x for x in zb))
    print("CPU half vs CPU half (same law by construction, R/2 each): mean z^2 = %.4f" % np.nanmean(zz**2))


if __name__ == "__main__":
    main()
