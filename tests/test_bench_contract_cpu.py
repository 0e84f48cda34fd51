"""bench.py's contract, the part that needs no GPU: the reference arm (`--impl reference`: the unmodified
reference on the host cores, the one place outside tests/ and smoke() where oracle/ may run) prints exactly ONE
line on stdout -- the JSON record with the keys the driver reads -- whatever libraries write there; and the
product arm refuses to run without a GPU instead of falling back to anything."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def run(*args):
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py")] + list(args), capture_output=True,
                          text=True, cwd=ROOT, timeout=600)


def test_reference_arm_prints_one_json_line():
    r = run("--impl", "reference", "--steps", "1", "--warmup", "0", "--cpu-sites", "1500", "--cpu-procs", "2")
    assert r.returncode == 0, r.stderr[-2000:]
    lines = r.stdout.splitlines()
    assert len(lines) == 1, r.stdout
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "mh_site_branch_updates_per_sec" and d["unit"] == "updates/s"
    assert d["higher_is_better"] is True and d["value"] > 0 and d["steps"] == 1 and d["n_gpus"] == 1
    assert d["cpu_baseline"]["kind"] in ("reference", "port") and d["cpu_baseline"]["cores"] == 2
    assert d["cpu_baseline"]["value"] == d["value"] and d["e2e"]["value"] == d["value"]
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0
    assert d["config"]["workload"] == "bal8" and d["gpu_launches"] == 0


def test_reference_arm_on_other_ranks_is_silent():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2"],
                       capture_output=True, text=True, cwd=ROOT, env=env, timeout=120)
    assert r.returncode == 0 and r.stdout == ""


def test_product_arm_fails_loudly_without_a_gpu():
    import torch
    if torch.cuda.is_available():
        import pytest
        pytest.skip("a GPU is present")
    r = run("--steps", "1", "--warmup", "0", "--sites", "3000", "--no-cpu-baseline")
    assert r.returncode != 0
    assert r.stdout.strip() == ""                 # no record, no fallback number
