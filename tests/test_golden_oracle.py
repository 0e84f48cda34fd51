"""Pins the CPU oracle (and the host-side file formats) against the committed golden
vectors in tests/golden/, which were produced by the unmodified reference build
(tests/golden/make_golden.py).  Runs anywhere -- /root/reference is not needed."""
import os

import numpy as np
import pytest

import cpu_libs
from epievo_b200.paths import PathSet, Tree, read_paths, read_tree, write_paths
from synth import rel_close

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CLI = os.path.join(GOLD, "cli")
CASES = ["testtree", "pair", "bal8", "bal32"]
SITES = (1, 2, 29, 58)


def load_case(name):
    g = np.load(os.path.join(GOLD, "lib_%s.npz" % name))
    tree = Tree.from_preorder(g["subtree_sizes"], g["branches"])
    ps = PathSet(int(g["n_sites"]), tree.n_nodes, g["init_state"], g["row_off"], g["jumps"],
                 g["tot_time"])
    return g, tree, ps


def fmt_params(T, baseline):
    """EpiEvoModel::format_for_param_file (EpiEvoModel.cpp:192-200), default ostream precision."""
    return "stationary\t%g\t%g\nbaseline\t%g\t%g\n" % (T[0], T[3], baseline[0], baseline[3])


@pytest.fixture(scope="module")
def o():
    return cpu_libs.oracle()


def test_ctmc_known_answers(o):
    g = np.load(os.path.join(GOLD, "lib_ctmc.npz"))
    for i in range(len(g["t"])):
        assert rel_close(o.trans_prob_mat(g["r0"][i], g["r1"][i], g["t"][i]), g["P"][i], 1e-13)
        assert rel_close(np.concatenate(o.expectations(g["r0"][i], g["r1"][i], g["t"][i])), g["E"][i], 1e-12)


@pytest.mark.parametrize("name", CASES)
def test_deterministic_passes(o, name):
    g, tree, ps = load_case(name)
    rates, T = g["rates"], g["T"]
    o.load_paths(ps)
    J, D = o.get_suff_stats()
    assert np.array_equal(J, g["J"]) and rel_close(D, g["D"], 1e-13)
    assert rel_close(o.tri_llh(rates, T), g["tri_llh"], 1e-13)
    for s in SITES:
        d = o.site_debug(tree, rates, T, s)
        for k, v in d.items():
            ref = g["dbg%d_%s" % (s, k)]
            if np.asarray(v).dtype.kind in "ui":
                assert np.array_equal(v, ref), (s, k)
            else:
                assert rel_close(v, ref, 1e-12, atol=1e-300), (s, k)
    o.scale_jump_times(g["scale_branches"])
    assert rel_close(o.store_paths().jumps, g["scaled_jumps"], 1e-15)


@pytest.mark.parametrize("name", CASES)
def test_mcmc_chain_and_m_step(o, name):
    """The reference's sequential sampler from mt19937 seed 4242: the C port must walk the
    same chain (same libstdc++ distributions re-implemented bit-compatibly)."""
    g, tree, ps = load_case(name)
    rates, T = g["rates"], g["T"]
    o.load_paths(ps)
    o.seed(4242)
    J, D, acc, _ = o.run_mcmc(tree, rates, T, 2, 3)
    assert np.array_equal(J, g["mcmc_J"]) and rel_close(D, g["mcmc_D"], 1e-12)
    assert acc == float(g["mcmc_acc"])
    out = o.store_paths()
    assert np.array_equal(out.row_off, g["mcmc_off"]) and np.array_equal(out.init_state, g["mcmc_init"])
    assert rel_close(out.jumps, g["mcmc_jumps"], 1e-12)
    llh, r2, T2, bl2 = o.estimate_rates(g["mcmc_J"], g["mcmc_D"], rates)
    assert rel_close(llh, g["m_llh"], 1e-12) and rel_close(r2, g["m_rates"], 1e-11)
    assert rel_close(T2, g["m_T"], 1e-11) and rel_close(bl2, g["m_baseline"], 1e-10, atol=1e-12)
    llh, r3, br3, T3, _ = o.estimate_rates_and_branches(g["mcmc_J"], g["mcmc_D"], rates, tree.branches)
    assert rel_close(llh, g["mb_llh"], 1e-12) and rel_close(r3, g["mb_rates"], 1e-11)
    assert rel_close(br3, g["mb_branches"], 1e-11) and rel_close(T3, g["mb_T"], 1e-11)


def read_param_file(path):
    rows = [l.split() for l in open(path)]
    return float(rows[0][1]), float(rows[0][2]), float(rows[1][1]), float(rows[1][2])


def test_cli_est_complete(o):
    """epievo_est_complete on the walk-through data: J/D pass + gradient ascent must give
    the parameter file the reference binary wrote (text-identical at its 6 digits)."""
    tree = read_tree(os.path.join(CLI, "tree.nwk"))
    ps = read_paths(os.path.join(CLI, "sim.paths"))
    assert ps.n_sites == 400 and ps.n_nodes == 5 and ps.node_names == tree.node_names
    rates, T, _, _ = o.model_from_params(*read_param_file(os.path.join(CLI, "test.param")))
    o.load_paths(ps)
    J, D = o.get_suff_stats()
    _, _, T2, bl2 = o.estimate_rates(J, D, rates)
    assert fmt_params(T2, bl2) == open(os.path.join(CLI, "est_complete.param")).read()


def replay_mcem(o, optimize_branches):
    tree = read_tree(os.path.join(CLI, "tree.nwk"))
    ps = read_paths(os.path.join(CLI, "sim.paths"))
    rates, T, _, bl = o.model_from_params(*read_param_file(os.path.join(CLI, "test.param")))
    o.load_paths(ps)
    o.seed(1)
    branches = np.array(tree.branches)
    log = []
    for itr in range(2):          # prog/epievo_est_params_histories.cpp:239-264, -i 2 -B 3 -L 2
        tree.branches = branches
        J, D, acc, _ = o.run_mcmc(tree, rates, T, 2, 3, reset=True)
        if optimize_branches:
            llh, rates, branches, T, bl = o.estimate_rates_and_branches(J, D, rates, branches)
        else:
            llh, rates, T, bl = o.estimate_rates(J, D, rates)
            rates, branches = o.set_one_change(rates, branches)
        o.scale_jump_times(branches)
        log.append("%d\t%g\t%g\t%g\t%g\t%g\t%g" % (itr + 1, T[0], T[3], bl[0], bl[3], acc, llh))
    return tree, T, bl, branches, log, o.store_paths()


def test_cli_mcem_replay(o):
    """Whole MCEM trajectory of epievo_est_params_histories -s 1 -i 2 -B 3 -L 2 replayed
    through the oracle: verbose log rows, final parameter file and the final local-paths
    file (byte-identical through our writer)."""
    tree, T, bl, _, log, out = replay_mcem(o, False)
    gold_log = [l.rstrip("\n") for l in open(os.path.join(CLI, "mcem.log"))]
    assert log == [l for l in gold_log if l[:2] in ("1\t", "2\t")]
    assert fmt_params(T, bl) == open(os.path.join(CLI, "mcem.param")).read()
    tmp = os.path.join(CLI, "_replay.paths")
    try:
        write_paths(tmp, out, tree.node_names)
        assert open(tmp).read() == open(os.path.join(CLI, "mcem.paths")).read()
    finally:
        if os.path.exists(tmp):
            os.remove(tmp)


def test_cli_mcem_branches_replay(o):
    tree, T, bl, branches, log, _ = replay_mcem(o, True)
    gold_log = [l.rstrip("\n") for l in open(os.path.join(CLI, "mcem_b.log"))]
    assert log == [l for l in gold_log if l[:2] in ("1\t", "2\t")]
    assert fmt_params(T, bl) == open(os.path.join(CLI, "mcem_b.param")).read()
    tree.branches = branches
    got = read_tree(os.path.join(CLI, "mcem_b.nwk"))
    assert rel_close(got.branches, branches, 1e-5)        # the tree file carries 6 digits


def test_paths_text_roundtrip():
    src = os.path.join(CLI, "mcem.paths")
    ps = read_paths(src)
    tmp = os.path.join(CLI, "_rt.paths")
    try:
        write_paths(tmp, ps)
        assert open(tmp).read() == open(src).read()
    finally:
        if os.path.exists(tmp):
            os.remove(tmp)
