#!/usr/bin/env python
"""Per-source-line instruction counts from an .ncu-rep captured with --import-source on
(compile with -lineinfo).  usage: ncu_lines.py <rep> [top_n]"""
import csv
import subprocess
import sys

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
filt = sys.argv[3] if len(sys.argv) > 3 else ""          # substring of the kernel name
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
rows = csv.reader(out.splitlines())
fname, hdr, agg, kernels, func = None, None, {}, 0, ""
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        fname = r[1].split("/")[-1]
        continue
    if r[0] == "Function Name":
        func = r[1]
        continue
    if r[0] == "Line No":
        hdr = r
        i_inst = hdr.index("Instructions Executed")
        i_thr = hdr.index("Thread Instructions Executed")
        i_smp = hdr.index("# Samples")
        continue
    if hdr and r[0] not in ("", "-") and filt in func:          # a source-line summary row
        try:
            key = (fname, int(r[0]), r[1].strip()[:90])
            a = agg.setdefault(key, [0, 0, 0])
            a[0] += int(r[i_inst]); a[1] += int(r[i_thr]); a[2] += int(r[i_smp])
        except (ValueError, IndexError):
            pass
tot = sum(a[0] for a in agg.values()) or 1
tots = sum(a[2] for a in agg.values()) or 1
print("total warp instructions (all captured launches): %d" % tot)
print("%6s %6s %5s  %s" % ("inst%", "smpl%", "lanes", "source"))
for key, a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print("%6.2f %6.2f %5.1f  %s:%d  %s" % (100.0*a[0]/tot, 100.0*a[2]/tots, a[1]/max(1, a[0]), key[0], key[1], key[2]))
