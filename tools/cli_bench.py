"""Wall time of the REAL drop-in: the reference's programs compiled unchanged against the B200 shim
(integration/_build/*_b200) next to the reference's own binaries (oracle/_ref/bin), same files, same flags.

  C1  test tree, 10^4 sites, `epievo_est_params_histories -s 1` with its defaults (10 iterations, B=L=10)
      input: `epievo_sim -s 1 -n 10000` -> `global_jumps_to_paths`, by the reference binaries (BASELINE config 1)
  C1M the same program on a 10^6-site local-paths file (histories from the library's simulator, written by
      epv_pathfile_write): reference `-i 1` (one iteration, ~2 minutes), drop-in `-i 1` and `-i 10`
  C3  `epievo_est_complete` (J/D pass over every path + M-step; BASELINE config 3's program) on the same 10^6-site
      file: the drop-in replaces get_sufficient_statistics / estimate_rates on paths at link time
  C2  `epievo_sim_pairwise -T 1.0 -L 10 -s 3` on 10^6 sites (root + leaf columns of `epievo_sim -s 2 -T 1.0`)

The drop-in's own phases come from EPIEVO_B200_TIMING=1 (flatten / upload / device / download / unflatten,
printed by host/SingleSiteSampler_b200.cpp); what is left of the wall time is the program's file I/O + M-step.
Writes one JSON object; run on a GPU box:  python tools/cli_bench.py > gpurun_out/r02_cli_bench.json
"""
import json
import os
import re
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
REF = os.path.join(ROOT, "oracle", "_ref", "bin")
B200 = os.path.join(ROOT, "integration", "_build")
CLI = os.path.join(ROOT, "tests", "golden", "cli")
TREE, PARAM = os.path.join(CLI, "tree.nwk"), os.path.join(CLI, "test.param")


def timed(cmd, env=None, timeout=3000):
    e = dict(os.environ)
    e.update(env or {})
    t0 = time.perf_counter()
    r = subprocess.run(cmd, capture_output=True, text=True, env=e, timeout=timeout)
    dt = time.perf_counter() - t0
    if r.returncode != 0:
        raise RuntimeError("%s failed:\n%s" % (" ".join(cmd), r.stderr[-2000:]))
    out = {"wall_s": round(dt, 3)}
    m = re.search(r"\[epievo_b200 timing\] (.*)", r.stderr)
    if m:
        ph = {k: float(v) for k, v in (kv.split("=") for kv in m.group(1).split())}
        out["phases_s"] = ph
        out["file_io_and_m_step_s"] = round(dt - sum(v for k, v in ph.items() if k != "calls"), 3)
    return out


def pair(name, args, tmp, gpus=(1,), ref=True, ref_args=None):
    res = {}
    if ref:
        res["reference"] = timed([os.path.join(REF, name)] + (ref_args or args) +
                                 ["-o", os.path.join(tmp, "ref.out")])
        res["reference"]["args"] = " ".join(a if not a.startswith("/") else os.path.basename(a) for a in (ref_args or args))
    for g in gpus:
        key = "b200" if g == 1 else "b200_%dgpu" % g
        res[key] = timed([os.path.join(B200, name + "_b200")] + args + ["-o", os.path.join(tmp, "b200.out")],
                         env={"EPIEVO_B200_TIMING": "1", "EPIEVO_B200_GPUS": str(g)})
        res[key]["args"] = " ".join(a if not a.startswith("/") else os.path.basename(a) for a in args)
    return res


def main():
    import numpy as np
    import torch

    from epievo_b200.device import Context
    from epievo_b200.paths import parse_newick, write_paths_native
    n_gpu = torch.cuda.device_count()
    out = {"gpus_present": n_gpu, "host_cores": os.cpu_count()}
    with tempfile.TemporaryDirectory() as tmp:
        # ---- C1: exactly the README walk-through at 10^4 sites --------------------------------------
        t0 = time.perf_counter()
        subprocess.run([REF + "/epievo_sim", "-s", "1", "-n", "10000", "-t", TREE, "-p", tmp + "/c1.jumps", PARAM,
                        tmp + "/c1.states"], check=True, capture_output=True)
        subprocess.run([REF + "/global_jumps_to_paths", "-t", TREE, tmp + "/c1.states", tmp + "/c1.jumps",
                        tmp + "/c1.paths"], check=True, capture_output=True)
        out["C1_input_by_reference_s"] = round(time.perf_counter() - t0, 2)
        c1 = ["-s", "1", "-p", tmp + "/c1.param", PARAM, TREE, tmp + "/c1.paths"]
        out["C1_est_params_histories_1e4_sites_10_iterations"] = pair("epievo_est_params_histories", c1, tmp)
        r = out["C1_est_params_histories_1e4_sites_10_iterations"]
        r["speedup"] = round(r["reference"]["wall_s"]/r["b200"]["wall_s"], 2)

        # ---- C1M: 10^6 sites, histories from the library's simulator --------------------------------
        tree = parse_newick(open(TREE).read())
        rates, T = np.zeros(8), np.zeros(4)
        from epievo_b200 import _lib as L
        d = lambda a: a.ctypes.data_as(L.dp)
        L.load().epv_model_from_params(0.844912, 0.893359, -0.8, -1.8, 1, d(rates), d(T), d(np.zeros(4)), d(np.zeros(4)))
        ctx = Context(0)
        ctx.set_tree(tree); ctx.set_model(rates, T)
        ps = ctx.simulate(1_000_000, 5, tree.branches, node_names=tree.node_names)
        ctx.close()
        t0 = time.perf_counter()
        write_paths_native(tmp + "/m.paths", ps, tree.node_names)
        out["C1M_paths_file_bytes"] = os.path.getsize(tmp + "/m.paths")
        out["C1M_write_paths_native_s"] = round(time.perf_counter() - t0, 2)
        base = ["-s", "1", "-p", tmp + "/m.param", PARAM, TREE, tmp + "/m.paths"]
        gp = tuple(g for g in (1, 2, 4, 8) if g <= n_gpu)
        res = pair("epievo_est_params_histories", ["-i", "1"] + base, tmp, gpus=gp)
        res["speedup_1_iteration"] = round(res["reference"]["wall_s"]/res["b200"]["wall_s"], 2)
        out["C1M_est_params_histories_1e6_sites_1_iteration"] = res
        out["C1M_est_params_histories_1e6_sites_10_iterations"] = pair("epievo_est_params_histories", base, tmp,
                                                                       gpus=gp, ref=False)
        # ---- C3-like: epievo_est_complete (J/D pass + M-step, no sampling) on the same 10^6-site file --
        c3 = [PARAM, TREE, tmp + "/m.paths"]
        res = pair("epievo_est_complete", c3, tmp)
        res["speedup"] = round(res["reference"]["wall_s"]/res["b200"]["wall_s"], 2)
        res["same_parameter_file"] = open(tmp + "/ref.out").read() == open(tmp + "/b200.out").read()
        out["C3_est_complete_1e6_sites"] = res
        del ps

        # ---- C2: pairwise, 10^6 sites, T = 1 ----------------------------------------------------------
        subprocess.run([REF + "/epievo_sim", "-s", "2", "-n", "1000000", "-T", "1.0", PARAM, tmp + "/p.states"],
                       check=True, capture_output=True)
        c2 = ["-T", "1.0", "-L", "10", "-s", "3", PARAM, tmp + "/p.states"]
        res = pair("epievo_sim_pairwise", c2, tmp)
        res["speedup"] = round(res["reference"]["wall_s"]/res["b200"]["wall_s"], 2)
        out["C2_sim_pairwise_1e6_sites_L10"] = res
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
