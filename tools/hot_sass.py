#!/usr/bin/env python
"""List the SASS instructions of a captured kernel that execute at least MIN times per env-step,
in program order (from `ncu --set full --import-source on`).
    python tools/hot_sass.py gpurun_out/prof_fused.ncu-rep ENV_STEPS_PER_LAUNCH [MIN=0.04]"""
import csv
import io
import subprocess
import sys

rep, steps = sys.argv[1], float(sys.argv[2])
thr = float(sys.argv[3]) if len(sys.argv) > 3 else 0.04
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
ins = []
for idx, r in enumerate(rows[2:]):
    try:
        ins.append((idx, r[1].strip(), int(r[5]) / steps, int(r[2])))
    except (ValueError, IndexError):
        pass
sst = sum(o[3] for o in ins) or 1
print("total warp instructions per env-step: %.1f" % sum(o[2] for o in ins))
for o in ins:
    if o[2] >= thr:
        print("%5d %6.3f %5.2f%%  %s" % (o[0], o[2], 100.0 * o[3] / sst, o[1]))
