#!/usr/bin/env python
"""Compile the CUDA library of the current tree as build/variants/NAME.so with -Xptxas -v and print, per kernel,
registers / spills / SASS instruction count (cuobjdump).   python tools/ptxas_info.py NAME [extra nvcc flags]"""
import os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
name, extra = sys.argv[1], sys.argv[2:]
out = os.path.join(ROOT, "build", "variants", name + ".so")
os.makedirs(os.path.dirname(out), exist_ok=True)
cmd = ["nvcc", "-std=c++17", "-O3", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-Xcompiler",
       "-fPIC,-fvisibility=hidden", "-shared", "-Xptxas", "-v", "-I" + ROOT + "/include",
       "-I" + ROOT + "/gym_eldorado_b200/csrc"] + extra + [ROOT + "/gym_eldorado_b200/csrc/cog_capi.cu", "-o", out]
r = subprocess.run(cmd, capture_output=True, text=True)
if r.returncode:
    print(r.stderr); sys.exit(1)
info, cur = {}, None
for l in r.stderr.splitlines():
    m = re.search(r"Compiling entry function '(\S+)'", l)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0]
    m = re.search(r"Used (\d+) registers", l)
    if m and cur: info.setdefault(cur, {})["regs"] = int(m.group(1))
    m = re.search(r"(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads", l)
    if m and cur: info.setdefault(cur, {}).update(stack=int(m.group(1)), spill=int(m.group(2)))
sass = subprocess.run(["cuobjdump", "-sass", out], capture_output=True, text=True).stdout
cur = None
for l in sass.splitlines():
    m = re.search(r"Function : (\S+)", l)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0]
        info.setdefault(cur, {})["sass"] = 0
    elif cur and re.match(r"\s+/\*[0-9a-f]{4,}\*/", l):
        info[cur]["sass"] += 1
        for op in ("BSSY", "BRA.DIV", "WARPSYNC"):
            if " " + op in l: info[cur][op] = info[cur].get(op, 0) + 1
for k, v in sorted(info.items()):
    print("%-48s regs %3s stack %4s spill %4s sass %6s BSSY %3s BRA.DIV %3s" % (k[:48], v.get("regs"), v.get("stack"), v.get("spill"), v.get("sass"), v.get("BSSY", 0), v.get("BRA.DIV", 0)))
