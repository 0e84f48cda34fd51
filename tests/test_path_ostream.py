"""host/path_ostream_b200.cpp replaces the reference's operator<<(ostream, Path) (Path.cpp:62-71) at link time:
the local-paths rows it prints must be byte-identical with the reference's.  Needs the reference sources and
the objects oracle/Makefile built from them (skipped elsewhere)."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = "/root/reference/src"
OBJ = os.path.join(ROOT, "oracle", "_ref", "obj")
NEW = os.path.join(ROOT, "integration", "_build", "path_ostream_b200.o")

pytestmark = pytest.mark.skipif(not (os.path.exists(REF + "/libepievo/Path.hpp") and os.path.exists(OBJ + "/Path.o")
                                     and os.path.exists(NEW)), reason="reference sources / oracle/_ref / integration/_build absent")


def test_replacement_operator_writes_the_same_bytes(tmp_path):
    src = os.path.join(ROOT, "tests", "cpp", "ostream_check.cpp")
    base = ["g++", "-std=c++11", "-O2", "-include", "limits", "-w", "-I" + REF + "/libepievo", "-I" + REF + "/smithlab_cpp", src]
    objs = [OBJ + "/Path.o", OBJ + "/sl_smithlab_utils.o", OBJ + "/sl_smithlab_os.o"]
    subprocess.run(base + objs + ["-o", str(tmp_path / "ref")], check=True)
    subprocess.run(base + [NEW] + objs + ["-Wl,--allow-multiple-definition", "-o", str(tmp_path / "new")], check=True)
    subprocess.run([str(tmp_path / "ref"), str(tmp_path / "a.txt")], check=True)
    subprocess.run([str(tmp_path / "new"), str(tmp_path / "b.txt")], check=True)
    a, b = open(tmp_path / "a.txt", "rb").read(), open(tmp_path / "b.txt", "rb").read()
    assert len(a) > 4_000_000 and a == b
