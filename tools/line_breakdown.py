#!/usr/bin/env python
"""Warp instructions per unit of work for every source line of one file, from an ncu capture taken with
--import-source on.    python tools/line_breakdown.py REPORT.ncu-rep UNITS_PER_LAUNCH cog_kernels.cuh [first last]"""
import csv
import io
import subprocess
import sys

rep, units, fname = sys.argv[1], float(sys.argv[2]), sys.argv[3]
lo, hi = (int(sys.argv[4]), int(sys.argv[5])) if len(sys.argv) > 5 else (0, 1 << 30)
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
cur, tot = None, 0
for r in csv.reader(io.StringIO(out)):
    if len(r) >= 2 and r[0] == "File Path":
        cur = r[1].split("/")[-1]
    elif len(r) > 8 and r[0].isdigit() and cur == fname:
        try:
            line, n = int(r[0]), int(r[7])
        except ValueError:
            continue
        if n and lo <= line <= hi:
            tot += n
            print("%5d %7.1f  %s" % (line, n / units, r[1].strip()[:120]))
print("total %.1f" % (tot / units))
