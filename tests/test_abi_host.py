"""CPU-side checks of the product library: it loads, exports every symbol the public
header declares, fails loudly without a GPU (no CPU fallback), and its host-side O(8)
model arithmetic / M-step matches the reference's golden results."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import cpu_libs
from epievo_b200 import _lib as L
from epievo_b200 import build as B
from synth import rel_close, TEST_PARAMS
from test_golden_oracle import CLI, fmt_params, load_case, read_param_file

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    B.build()
    return L.load()


def header_symbols():
    text = open(os.path.join(ROOT, "include", "epievo_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(epv_[a-z0-9_]+)\s*\(", text)))


def test_every_declared_symbol_is_exported_and_bound(lib):
    syms = header_symbols()
    assert len(syms) >= 30
    for s in syms:
        assert hasattr(lib, s), "missing export: " + s
        assert s in L.SIGNATURES, "ctypes binding missing for " + s
    assert sorted(L.SIGNATURES) == syms


def test_library_is_sm100a_native():
    out = os.popen("cuobjdump -lelf %s 2>/dev/null" % L.LIB_PATH).read()
    if not out:
        pytest.skip("cuobjdump not available")
    assert "sm_100a" in out


def test_no_cpu_fallback(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    h = L.ctx_p()
    assert lib.epv_create(C.byref(h), 0) != 0 and not h.value
    from epievo_b200.device import Context, EpvError
    with pytest.raises(EpvError):
        Context(0)


def _d(a):
    return a.ctypes.data_as(L.dp)


def test_model_arithmetic_matches_reference(lib):
    o = cpu_libs.oracle()
    for params in (TEST_PARAMS, (0.7, 0.6, -0.3, -1.1), (0.95, 0.9, -2.0, -0.5)):
        for scale in (0, 1):
            r, T, Q, bl = np.zeros(8), np.zeros(4), np.zeros(4), np.zeros(4)
            lib.epv_model_from_params(*params, scale, _d(r), _d(T), _d(Q), _d(bl))
            for got, want in zip((r, T, Q, bl), o.model_from_params(*params, scale)):
                assert rel_close(got, want, 1e-12, atol=1e-15)
        rates = o.model_from_params(*params, 1)[0]
        assert rel_close(lib.epv_rate_scaling_factor(_d(rates)), o.rate_scaling_factor(rates), 1e-12)
        T, Q, bl = np.zeros(4), np.zeros(4), np.zeros(4)
        lib.epv_model_rebuild(_d(rates), _d(T), _d(Q), _d(bl))
        for got, want in zip((T, Q, bl), o.model_rebuild(rates)):
            assert rel_close(got, want, 1e-11, atol=1e-13)


@pytest.mark.parametrize("name", ["testtree", "pair", "bal8", "bal32"])
def test_m_step_matches_golden(lib, name):
    g, tree, _ = load_case(name)
    J = np.ascontiguousarray(g["mcmc_J"]).reshape(-1)
    D = np.ascontiguousarray(g["mcmc_D"]).reshape(-1)
    r, T, bl = g["rates"].copy(), np.zeros(4), np.zeros(4)
    llh = lib.epv_estimate_rates(tree.n_nodes, _d(J), _d(D), 1e-10, _d(r), _d(T), _d(bl))
    assert rel_close(llh, g["m_llh"], 1e-9) and rel_close(r, g["m_rates"], 1e-9)
    assert rel_close(T, g["m_T"], 1e-9) and rel_close(bl, g["m_baseline"], 1e-9, atol=1e-12)
    r, br = g["rates"].copy(), np.array(tree.branches)
    llh = lib.epv_estimate_rates_and_branches(tree.n_nodes, _d(J), _d(D), 1e-10, _d(r), _d(br), _d(T), _d(bl))
    assert rel_close(llh, g["mb_llh"], 1e-9) and rel_close(r, g["mb_rates"], 1e-9)
    assert rel_close(br, g["mb_branches"], 1e-9) and rel_close(T, g["mb_T"], 1e-9)
    o = cpu_libs.oracle()
    r, br = g["rates"].copy(), np.array(tree.branches)
    lib.epv_set_one_change(_d(r), tree.n_nodes, _d(br))
    wr, wb = o.set_one_change(g["rates"], tree.branches)
    assert rel_close(r, wr, 1e-12) and rel_close(br, wb, 1e-12)


def test_est_complete_param_file_from_product_m_step(lib):
    """Golden J/D of the CLI walk-through -> product M-step -> the reference's param file."""
    from epievo_b200.paths import read_paths
    o = cpu_libs.oracle()
    ps = read_paths(os.path.join(CLI, "sim.paths"))
    rates = o.model_from_params(*read_param_file(os.path.join(CLI, "test.param")))[0]
    o.load_paths(ps)
    J, D = o.get_suff_stats()
    r, T, bl = rates.copy(), np.zeros(4), np.zeros(4)
    lib.epv_estimate_rates(ps.n_nodes, _d(J.reshape(-1).copy()), _d(D.reshape(-1).copy()), 1e-10,
                           _d(r), _d(T), _d(bl))
    assert fmt_params(T, bl) == open(os.path.join(CLI, "est_complete.param")).read()


def test_multi_partition_matches_the_python_driver():
    """epv_multi_partition (csrc/api_multi.cu, the C++ drop-in's sharding) == epievo_b200.sharded.partition:
    contiguous blocks that cover the sites, starts multiples of 3, at least 6 sites each."""
    import ctypes as C

    import numpy as np

    from epievo_b200 import _lib as L
    from epievo_b200 import sharded as S
    lib = L.load()
    for n, w in ((100, 3), (1001, 8), (12, 2), (10**8, 8), (3*10**7, 4), (47, 7)):
        b, e = np.zeros(w, np.uint64), np.zeros(w, np.uint64)
        rc = lib.epv_multi_partition(n, w, b.ctypes.data_as(L.u64p), e.ctypes.data_as(L.u64p))
        try:
            want = S.partition(n, w)
        except ValueError:
            assert rc != 0
            continue
        assert rc == 0
        assert [(int(x), int(y)) for x, y in zip(b, e)] == want
        assert b[0] == 0 and e[-1] == n and all(int(x) % 3 == 0 for x in b)
    b, e = np.zeros(4, np.uint64), np.zeros(4, np.uint64)
    assert lib.epv_multi_partition(20, 4, b.ctypes.data_as(L.u64p), e.ctypes.data_as(L.u64p)) != 0   # < 6 sites per shard


@pytest.mark.parametrize("exe,args", [
    ("epievo_est_params_histories_b200", ["-s", "1", "-i", "1", "-o", "{tmp}/o.paths", "-p", "{tmp}/o.param",
                                          "{cli}/test.param", "{cli}/tree.nwk", "{cli}/sim.paths"]),
    ("epievo_est_complete_b200", ["-o", "{tmp}/c.param", "{cli}/test.param", "{cli}/tree.nwk", "{cli}/sim.paths"]),
])
def test_dropin_programs_fail_the_reference_way_without_a_gpu(tmp_path, exe, args):
    """The reference's own main()s, compiled unchanged against the drop-in (integration/Makefile): without a
    CUDA device they print the library's message and return EXIT_FAILURE through the reference's own
    `catch (const std::exception &)` -- they neither crash nor compute anything on the host."""
    import subprocess

    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    p = os.path.join(ROOT, "integration", "_build", exe)
    if not os.path.exists(p):
        pytest.skip("integration/_build not built (make -C integration needs /root/reference)")
    r = subprocess.run([p] + [a.format(tmp=tmp_path, cli=CLI) for a in args], capture_output=True, text=True, timeout=120)
    assert r.returncode == 1, (r.returncode, r.stderr[-500:])
    assert "no usable CUDA device" in r.stderr and "no CPU fallback" in r.stderr
    assert not os.path.exists(os.path.join(str(tmp_path), "c.param")) or os.path.getsize(os.path.join(str(tmp_path), "c.param")) == 0
