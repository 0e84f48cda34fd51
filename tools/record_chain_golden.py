"""Records the chains of the CURRENT library on fixed inputs into tests/golden/chain_golden.npz
(run on a GPU box; committed).  tests/test_gpu_chain_golden.py then holds later kernels to them:
same Philox streams + same arithmetic => the same histories bit for bit.  Recorded with the round-1
update kernel (v8) before the lockstep rewrite of round 2."""
import hashlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np

CASES = [  # name, tree, n_sites, generator, density, burn, batch
    ("testtree", "testtree", 3000, "synth", 1.0),
    ("testtree_dense", "testtree", 600, "synth", 12.0),
    ("bal8", "bal8", 3000, "sim", 0.0),
    ("bal8_dense", "bal8", 900, "synth", 6.0),
    ("pair", "pair", 2000, "sim", 0.0),
    ("bal32", "bal32", 400, "sim", 0.0),
    ("bal64", "bal64", 200, "sim", 0.0),
]


def make_case(name, tree_name, n, gen, density):
    import cpu_libs
    from epievo_b200.paths import Tree
    from epievo_b200.synthetic import simulate_paths
    from synth import TEST_PARAMS, ref_test_tree, synth_paths
    rates, T, _, _ = cpu_libs.oracle().model_from_params(*TEST_PARAMS)
    tree = {"testtree": ref_test_tree, "pair": lambda: Tree.pair(1.0), "bal8": lambda: Tree.balanced(8, 0.05),
            "bal32": lambda: Tree.balanced(32, 0.05), "bal64": lambda: Tree.balanced(64, 0.03)}[tree_name]()
    ps = synth_paths(tree, n, 5, density) if gen == "synth" else simulate_paths(tree, n, rates, T, seed=77)
    return tree, rates, T, ps


def run_case(ctx, tree, rates, T, ps):
    ctx.set_tree(tree); ctx.set_model(rates, T); ctx.upload_paths(ps); ctx.reset()
    J, D, acc = ctx.run_mcmc(3, 3, seed=31, first_sweep=2)
    out = ctx.download_paths()
    h = hashlib.sha256()
    for a in (out.init_state, out.row_off, out.jumps):
        h.update(np.ascontiguousarray(a).tobytes())
    return dict(J=J, D=D, acc=np.float64(acc), tri=ctx.tri_llh(), n_jumps=np.int64(out.jumps.size),
                sha=np.frombuffer(h.digest(), np.uint8).copy())


if __name__ == "__main__":
    from epievo_b200.device import Context
    store = {}
    for name, tree_name, n, gen, density in CASES:
        tree, rates, T, ps = make_case(name, tree_name, n, gen, density)
        c = Context(0)
        for k, v in run_case(c, tree, rates, T, ps).items():
            store[name + "." + k] = v
        c.close()
        print(name, "acc", float(store[name + ".acc"]), "jumps", int(store[name + ".n_jumps"]))
    dst = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "chain_golden.npz")
    os.makedirs(os.path.dirname(dst), exist_ok=True)
    np.savez_compressed(dst, **store)
    print("wrote", dst)
