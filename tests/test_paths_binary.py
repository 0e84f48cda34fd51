"""Binary sidecar of the local-paths file (epv_pathfile_write_binary; SURVEY 8f-1's "optional binary CSR
sidecar"): same content as the text form, read through the same reader; damaged files are refused; the
read_paths drop-in prefers an up-to-date sidecar and falls back to the text.  Host code only."""
import os
import subprocess
import time

import numpy as np
import pytest

from epievo_b200.paths import PathSet, Tree, read_paths_native, write_paths_binary, write_paths_native
from epievo_b200.synthetic import simulate_paths
from synth import ref_test_tree

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CLI = os.path.join(ROOT, "tests", "golden", "cli")


def model():
    import bench
    return bench.test_model()


@pytest.mark.parametrize("tree,n,threads", [(ref_test_tree(), 3000, 1), (Tree.balanced(8, 0.05), 20011, 0),
                                            (Tree.pair(4.0), 5000, 3)])
def test_round_trip_equals_the_text_form(tmp_path, tree, n, threads):
    rates, T = model()
    ps = simulate_paths(tree, n, rates, T, seed=4)
    names = ["n%d" % i for i in range(tree.n_nodes)]
    txt, binf = str(tmp_path / "a.paths"), str(tmp_path / "a.epvb")
    write_paths_native(txt, ps, names)
    write_paths_binary(binf, ps, names, n_threads=threads)
    a, b = read_paths_native(txt), read_paths_native(binf, n_threads=threads)
    assert b.same_as(a) and b.node_names == a.node_names == names
    assert np.array_equal(b.tot_time, a.tot_time) and np.array_equal(b.row_off, a.row_off)
    # 2 bytes per row + 8 per jump + a small header: far smaller than the text
    rows = tree.n_nodes * n
    assert os.path.getsize(binf) <= 48 + 8 * tree.n_nodes + 8 * sum(len(x) + 1 for x in names) + 2 * rows + 8 + 8 * len(ps.jumps)
    assert os.path.getsize(binf) < 0.5 * os.path.getsize(txt)           # (8 vs ~19 bytes per jump, 2 vs ~22 per row)


def test_golden_text_file_through_the_sidecar(tmp_path):
    a = read_paths_native(os.path.join(CLI, "sim.paths"))
    f = str(tmp_path / "sim.epvb")
    write_paths_binary(f, a, a.node_names)
    b = read_paths_native(f)
    assert b.same_as(a) and b.node_names == a.node_names
    out = str(tmp_path / "back.paths")
    write_paths_native(out, b, b.node_names)
    assert open(out, "rb").read() == open(os.path.join(CLI, "sim.paths"), "rb").read()      # text -> binary -> text


def test_damaged_files_are_refused(tmp_path):
    rates, T = model()
    ps = simulate_paths(ref_test_tree(), 2000, rates, T, seed=1)
    f = str(tmp_path / "a.epvb")
    write_paths_binary(f, ps, ["r", "a", "b", "c", "d"])
    good = open(f, "rb").read()
    for name, data, msg in (("flip", good[:len(good) // 2] + bytes([good[len(good) // 2] ^ 1]) + good[len(good) // 2 + 1:], "checksum"),
                            ("cut", good[:-8], "truncated"),
                            ("header", good[:8] + (1 << 50).to_bytes(8, "little") + good[16:], "bad header"),
                            ("short", good[:20], "truncated")):
        p = str(tmp_path / (name + ".epvb"))
        open(p, "wb").write(data)
        with pytest.raises(RuntimeError, match=msg):
            read_paths_native(p)
    assert read_paths_native(f).same_as(ps)                 # the intact file still reads


def test_rows_beyond_the_format_are_an_error(tmp_path):
    n, n_nodes = 4, 2
    init = np.zeros(n * n_nodes, np.uint8)
    off = np.zeros(n * n_nodes + 1, np.uint64)
    off[n + 2:] = 40000                                     # one row of 40000 jumps on the branch
    ps = PathSet(n, n_nodes, init, off, np.sort(np.random.default_rng(0).random(40000)), np.array([0.0, 1.0]), ["r", "l"])
    with pytest.raises(RuntimeError, match="32767"):
        write_paths_binary(str(tmp_path / "x.epvb"), ps, ["r", "l"])
    assert not os.path.exists(str(tmp_path / "x.epvb"))


REF = "/root/reference/src"
OBJ = os.path.join(ROOT, "oracle", "_ref", "obj")
NEW = os.path.join(ROOT, "integration", "_build", "read_paths_b200.o")


@pytest.mark.skipif(not (os.path.exists(REF + "/libepievo/Path.hpp") and os.path.exists(OBJ + "/Path.o") and os.path.exists(NEW)),
                    reason="reference sources / oracle/_ref / integration/_build absent")
def test_read_paths_dropin_prefers_an_up_to_date_sidecar(tmp_path):
    """The reference's read_paths and the drop-in (text; sidecar; stale sidecar; damaged sidecar) return the same
    [node][site] Path matrix -- compared through a digest printed by tests/cpp/read_paths_check.cpp."""
    src = os.path.join(ROOT, "tests", "cpp", "read_paths_check.cpp")
    base = ["g++", "-std=c++11", "-O2", "-include", "limits", "-w", "-I" + REF + "/libepievo", "-I" + REF + "/smithlab_cpp", src]
    objs = [OBJ + "/Path.o", OBJ + "/sl_smithlab_utils.o", OBJ + "/sl_smithlab_os.o"]
    subprocess.run(base + objs + ["-o", str(tmp_path / "ref")], check=True)
    subprocess.run(base + [NEW] + objs + ["-L" + os.path.join(ROOT, "epievo_b200"), "-lepievo_b200",
                                          "-Wl,-rpath," + os.path.join(ROOT, "epievo_b200"), "-Wl,--allow-multiple-definition",
                                          "-pthread", "-o", str(tmp_path / "new")], check=True)
    rates, T = model()
    ps = simulate_paths(ref_test_tree(), 4000, rates, T, seed=8)
    names = ["G", "E", "C", "D", "F"]
    txt = str(tmp_path / "a.paths")
    write_paths_native(txt, ps, names)
    run = lambda exe, f: subprocess.run([str(tmp_path / exe), f], capture_output=True, text=True)
    want = run("ref", txt)
    assert want.returncode == 0 and "digest" in want.stdout
    assert run("new", txt).stdout == want.stdout                       # text, no sidecar
    write_paths_binary(txt + ".epvb", ps, names)
    assert run("new", txt).stdout == want.stdout                       # sidecar
    # a sidecar with OTHER content that is up to date is what gets read: proves the preference
    other = simulate_paths(ref_test_tree(), 4000, rates, T, seed=9)
    write_paths_binary(txt + ".epvb", other, names)
    assert run("new", txt).stdout != want.stdout
    # ... unless the text file is newer, or the sidecar is damaged
    time.sleep(0.05)
    os.utime(txt, None)
    assert run("new", txt).stdout == want.stdout
    write_paths_binary(txt + ".epvb", other, names)
    data = bytearray(open(txt + ".epvb", "rb").read()); data[-3] ^= 0x40
    open(txt + ".epvb", "wb").write(bytes(data))
    assert run("new", txt).stdout == want.stdout
