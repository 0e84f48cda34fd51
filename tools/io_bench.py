#!/usr/bin/env python
"""Local-paths text I/O: the library's multi-threaded reader / writer (epv_pathfile_*)
timed beside the reference's read_paths / writer statements (oracle/_ref) on the same file.
usage: io_bench.py [n_sites] [threads]   (CPU only; prints one JSON line)"""
import json
import os
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]

import cpu_libs  # noqa: E402
from epievo_b200.paths import read_paths_native, write_paths_binary, write_paths_native  # noqa: E402
from epievo_b200.synthetic import simulate_paths  # noqa: E402
from synth import TEST_PARAMS, ref_test_tree  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
threads = int(sys.argv[2]) if len(sys.argv) > 2 else 0
rates, T, _, _ = cpu_libs.oracle().model_from_params(*TEST_PARAMS)
tree = ref_test_tree()
ps = simulate_paths(tree, n, rates, T, seed=1)
ps.node_names = list(tree.node_names)
d = tempfile.mkdtemp()
f, g = os.path.join(d, "a.paths"), os.path.join(d, "b.paths")

t0 = time.perf_counter(); write_paths_native(f, ps, n_threads=threads); tw = time.perf_counter() - t0
size = os.path.getsize(f)
t0 = time.perf_counter(); back = read_paths_native(f, threads); tr = time.perf_counter() - t0
assert back.same_as(ps)
out = {"n_sites": n, "n_branches": ps.n_nodes - 1, "file_mb": size / 1e6,
       "threads": threads or os.cpu_count(),
       "native_write_s": tw, "native_read_s": tr,
       "native_write_mb_s": size / 1e6 / tw, "native_read_mb_s": size / 1e6 / tr}
# the binary sidecar of the same content
b = os.path.join(d, "a.epvb")
t0 = time.perf_counter(); write_paths_binary(b, ps, n_threads=threads); tbw = time.perf_counter() - t0
t0 = time.perf_counter(); back = read_paths_native(b, threads); tbr = time.perf_counter() - t0
assert back.same_as(ps)
out.update({"binary_mb": os.path.getsize(b) / 1e6, "binary_write_s": tbw, "binary_read_s": tbr})
os.remove(b)
if cpu_libs.ref_available():
    ref = cpu_libs.ref()
    got, sec = ref.read_paths_file(f, timing=True)
    assert got.same_as(ps)
    wsec = ref.write_paths_file(g, ps.node_names)
    assert open(f, "rb").read() == open(g, "rb").read()
    out.update({"reference_read_s": sec, "reference_write_s": wsec,
                "reference_read_mb_s": size / 1e6 / sec, "reference_write_mb_s": size / 1e6 / wsec,
                "read_speedup": sec / tr, "write_speedup": wsec / tw})
for p in (f, g):
    if os.path.exists(p):
        os.remove(p)
os.rmdir(d)
print(json.dumps(out))
