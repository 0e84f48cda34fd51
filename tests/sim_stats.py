"""Per-site statistics of a simulated data set, shared by tests/golden/make_sim_golden.py (the
reference's epievo_sim) and tests/test_gpu_sim.py (epv_simulate)."""
import numpy as np

import cpu_libs


def stats_of(ps):
    o = cpu_libs.oracle()
    o.load_paths(ps)
    J, D = o.get_suff_stats()                               # [n_nodes, 8], interior sites
    n_int = ps.n_sites - 2
    lens = ps.row_len()                                     # [n_nodes, n_sites]
    init = ps.init_state.reshape(ps.n_nodes, ps.n_sites).astype(np.int64)
    end = init ^ (lens & 1)
    return {"J": (J / n_int).tolist(), "D": (D / n_int).tolist(),
            "jumps_per_site": (lens[:, 1:-1].sum(1) / n_int).tolist(),
            "frac1_start": init[:, 1:-1].mean(1).tolist(), "frac1_end": end[:, 1:-1].mean(1).tolist()}
