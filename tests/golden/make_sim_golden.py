#!/usr/bin/env python
"""Statistics of the UNMODIFIED reference forward simulator (oracle/_ref/bin/epievo_sim +
global_jumps_to_paths, built by oracle/Makefile from /root/reference), for the parity test of
epv_simulate (tests/test_gpu_sim.py).  Run in the build container only; the output
tests/golden/sim_stats.json is committed.

For each case the reference binaries are run with seeds 1..R on N sites; per replicate the
per-branch triplet statistics J, D (add_sufficient_statistics on the simulated histories, through
the C oracle), the jump count per branch and the fraction of 1s at every node are taken per site;
the file holds their mean and standard error over the replicates.  The simulators share a law, not
a random stream, so this is what can be compared.
"""
import json
import os
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
for p in (ROOT, os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)

from epievo_b200.paths import read_paths_native as read_paths   # noqa: E402
from sim_stats import stats_of                             # noqa: E402

BIN = os.path.join(ROOT, "oracle", "_ref", "bin")
CLI = os.path.join(HERE, "cli")                             # tree.nwk, test.param (copies of the reference's)


def run_case(name, sim_args, conv_args, n_sites, reps):
    acc = {}
    with tempfile.TemporaryDirectory() as tmp:
        for seed in range(1, reps + 1):
            jumps, states, paths = (os.path.join(tmp, f) for f in ("j", "s", "p"))
            subprocess.run([BIN + "/epievo_sim", "-s", str(seed), "-n", str(n_sites)] + sim_args +
                           ["-p", jumps, CLI + "/test.param", states], check=True, capture_output=True)
            subprocess.run([BIN + "/global_jumps_to_paths"] + conv_args + [states, jumps, paths],
                           check=True, capture_output=True)
            st = stats_of(read_paths(paths))
            for k, v in st.items():
                acc.setdefault(k, []).append(np.asarray(v, np.float64))
    out = {"n_sites": n_sites, "replicates": reps}
    for k, v in acc.items():
        a = np.stack(v)
        out[k] = {"mean": a.mean(0).tolist(), "sem": (a.std(0, ddof=1) / np.sqrt(reps)).tolist()}
    print(name, "jumps/site per branch", np.round(out["jumps_per_site"]["mean"], 4))
    return out


def main():
    tree = CLI + "/tree.nwk"
    res = {
        "testtree": run_case("testtree", ["-t", tree], ["-t", tree], 50000, 60),
        "pair_T1": run_case("pair_T1", ["-T", "1.0"], ["-T", "1.0"], 50000, 60),
    }
    json.dump(res, open(os.path.join(HERE, "sim_stats.json"), "w"), indent=0)


if __name__ == "__main__":
    main()
