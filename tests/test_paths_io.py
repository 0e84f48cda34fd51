"""Local-paths text I/O: the library's multi-threaded reader / writer
(include/epievo_b200.h epv_pathfile_*) against the reference's own files and code.
* golden files written by the reference BINARIES (tests/golden/cli/*.paths): parsing gives
  what the line-by-line Python restatement of read_paths (Path.cpp:123-148) gives, and
  writing the parsed paths back reproduces the file byte for byte;
* the unmodified reference's read_paths (oracle/_ref, when built) reads files written by
  epv_pathfile_write into the same paths;
* any thread count, files without trailing newline, empty rows, long rows, error cases."""
import os

import numpy as np
import pytest

from epievo_b200.paths import (PathSet, Tree, read_paths, read_paths_native, write_paths,
                               write_paths_native)
from epievo_b200.synthetic import simulate_paths
from synth import ref_test_tree

GOLD = os.path.join(os.path.dirname(__file__), "golden", "cli")
FILES = [os.path.join(GOLD, f) for f in ("sim.paths", "mcem.paths")]


@pytest.mark.parametrize("path", FILES)
@pytest.mark.parametrize("threads", [1, 2, 3, 7, 0])
def test_golden_file_round_trip(tmp_path, path, threads):
    want = read_paths(path)
    got = read_paths_native(path, threads)
    assert got.same_as(want) and got.node_names == want.node_names
    out = str(tmp_path / "out.paths")
    write_paths_native(out, got, n_threads=threads)
    assert open(out, "rb").read() == open(path, "rb").read()


def _random_paths(seed, n_sites, tree):
    rng = np.random.default_rng(seed)
    rates = np.array([0.3, 1.1, 2.0, 0.7, 0.9, 3.0, 0.2, 1.5])
    T = np.array([0.8, 0.2, 0.3, 0.7])
    ps = simulate_paths(tree, n_sites, rates, T, seed=seed)
    ps.node_names = ["n%d" % i for i in range(ps.n_nodes)]
    # awkward doubles: subnormal-ish, tiny exponents, values needing all 17 digits
    if ps.jumps.size > 4:
        ps.jumps[:4] = np.sort(np.array([5e-324, 1.2345678901234567e-300, 1e-5, 0.1 + 0.2 - 0.3 + 1e-3]))
    return ps, rng


@pytest.mark.parametrize("n_sites,threads", [(1, 1), (2, 4), (999, 3), (20000, 0)])
def test_native_equals_python_writer_and_reader(tmp_path, n_sites, threads):
    ps, _ = _random_paths(5, max(n_sites, 3), ref_test_tree())
    if n_sites < 3:
        ps = ps.site_slice(0, n_sites)
        ps.node_names = ["n%d" % i for i in range(ps.n_nodes)]
    a, b = str(tmp_path / "a.paths"), str(tmp_path / "b.paths")
    write_paths(a, ps)
    write_paths_native(b, ps, n_threads=threads)
    assert open(a, "rb").read() == open(b, "rb").read()
    back = read_paths_native(b, threads)
    assert back.same_as(ps) and back.node_names == ps.node_names


def test_reference_reads_what_we_write(tmp_path):
    import cpu_libs
    if not cpu_libs.ref_available():
        pytest.skip("oracle/_ref not built")
    ref = cpu_libs.ref()
    ps, _ = _random_paths(9, 500, Tree.balanced(8, 0.05))
    f = str(tmp_path / "w.paths")
    write_paths_native(f, ps, n_threads=3)
    got = ref.read_paths_file(f)
    assert got.same_as(ps)
    # and the reference's own writer produces the same bytes as ours
    g = str(tmp_path / "r.paths")
    ref.write_paths_file(g, ps.node_names)
    assert open(g, "rb").read() == open(f, "rb").read()


def test_no_trailing_newline_and_crlf(tmp_path):
    ps, _ = _random_paths(2, 50, ref_test_tree())
    f = str(tmp_path / "x.paths")
    write_paths_native(f, ps, n_threads=1)
    raw = open(f, "rb").read()
    g = str(tmp_path / "y.paths")
    open(g, "wb").write(raw.rstrip(b"\n"))
    assert read_paths_native(g, 2).same_as(ps)


def test_long_row_and_empty_rows(tmp_path):
    tree = Tree.pair(1.0)
    n = 10
    init = np.zeros(2 * n, np.uint8)
    cnt = np.zeros(2 * n, np.int64)
    cnt[n + 3] = 5000
    off = np.zeros(2 * n + 1, np.uint64)
    off[1:] = np.cumsum(cnt)
    jumps = np.sort(np.random.default_rng(0).random(5000))
    ps = PathSet(n, 2, init, off, jumps, np.array([0.0, 1.0]), ["R", "L"])
    f = str(tmp_path / "l.paths")
    write_paths_native(f, ps, n_threads=4)
    assert read_paths_native(f, 4).same_as(ps)
    assert read_paths(f).same_as(ps)


def test_errors(tmp_path):
    with pytest.raises(RuntimeError, match="cannot read"):
        read_paths_native(str(tmp_path / "missing.paths"))
    bad = str(tmp_path / "bad.paths")
    open(bad, "w").write("NODE:a\nNODE:b\n0\t0\t0.5\t\nNODE:c\n0\t0\t0.5\t\n1\t0\t0.5\t\n")
    with pytest.raises(RuntimeError, match="ragged"):
        read_paths_native(bad)
    open(bad, "w").write("hello\n")
    with pytest.raises(RuntimeError, match="not a local-paths file"):
        read_paths_native(bad)
    open(bad, "w").write("NODE:a\nNODE:b\n0\tx\t0.5\t\n")
    with pytest.raises(RuntimeError, match="malformed"):
        read_paths_native(bad)
