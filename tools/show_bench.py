"""Prints the key figures of bench.py output files (the JSON line is the last line starting with '{'; NCCL may
print its version line to stdout before it).  usage: show_bench.py file..."""
import json
import sys

for f in sys.argv[1:]:
    try:
        line = [l for l in open(f) if l.startswith("{")][-1]
        d = json.loads(line)
    except Exception as e:
        print(f, "unreadable:", e)
        continue
    r = d.get("roofline") or {}
    e2e = d.get("e2e") or {}
    mc = d.get("mcem_iter") or {}
    print("%s: n_gpus %s %s value %.4g ms/step %.1f e2e %s upd %.3f reb %.3f jd %.3f mcem %s acc %.4f" % (
        f.split("/")[-1], d.get("n_gpus"), d.get("scaling"), d["value"], d["ms_per_step"],
        "%.4g" % e2e["value"] if e2e else None, r.get("avg_launch_ms", 0), r.get("rebuild_kernel_avg_ms", 0),
        (d.get("suffstats_pass") or {}).get("ms", 0), ("%.1f ms (%s GB)" % (mc["ms"], round(mc.get("peak_device_memory_gb", 0), 1))) if mc else None,
        d.get("acc_rate", 0)))
