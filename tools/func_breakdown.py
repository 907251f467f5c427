#!/usr/bin/env python
"""Warp instructions per unit of work by source function, from an ncu capture with --import-source on.
    python tools/func_breakdown.py REPORT.ncu-rep UNITS_PER_LAUNCH"""
import bisect
import csv
import io
import os
import re
import subprocess
import sys

rep, units = sys.argv[1], float(sys.argv[2])
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
cur, agg = None, []
for r in csv.reader(io.StringIO(out)):
    if len(r) >= 2 and r[0] == "File Path":
        cur = r[1].split("/")[-1]
    elif len(r) > 8 and r[0].isdigit():
        try:
            agg.append((cur, int(r[0]), int(r[7])))
        except ValueError:
            pass
funcs = {}
for name in ("cog_device.cuh", "cog_kernels.cuh"):
    fl = []
    src_dir = os.environ.get("COG_SRC_DIR", os.path.join(root, "gym_eldorado_b200", "csrc"))  # the sources the capture was built from
    for i, l in enumerate(open(os.path.join(src_dir, name)), 1):
        if l.startswith("__device__") or l.startswith("__global__") or l.startswith("template <"):
            if l.startswith("template <"):
                continue
            m = re.search(r"(\w+)\(", l)
            if m:
                fl.append((i, m.group(1)))
    funcs[name] = fl
tot = {}
for f, ln, c in agg:
    name = f
    if f in funcs:
        fl = funcs[f]
        k = bisect.bisect_right([x[0] for x in fl], ln) - 1
        name = fl[k][1] if k >= 0 else f
    tot[name] = tot.get(name, 0) + c
for k, v in sorted(tot.items(), key=lambda x: -x[1])[:30]:
    print("%8.1f  %s" % (v / units, k))
print("%8.1f  attributed total" % (sum(tot.values()) / units))
