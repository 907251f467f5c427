#!/usr/bin/env python
"""Per kernel of a built library: SASS instructions and local-memory stores / loads of the kernel body, split from an
out-of-line callee at the call target, and how many of them sit next to the call (cuobjdump).   python tools/sass_call_stats.py lib.so"""
import subprocess,sys,re
so=sys.argv[1]
sass=subprocess.run(["cuobjdump","-sass",so],capture_output=True,text=True).stdout
cur=None; funcs={}
for l in sass.splitlines():
    m=re.search(r"Function : (\S+)",l)
    if m:
        cur=subprocess.run(["c++filt",m.group(1)],capture_output=True,text=True).stdout.strip().split("(")[0]; funcs[cur]=[]; continue
    m=re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);",l)
    if m and cur: funcs[cur].append((int(m.group(1),16),m.group(2)))
for f,ins in funcs.items():
    if not any(k in f for k in ("k_reset","k_step_tail<false, false>","k_step_single<true>")): continue
    calls=[(i,x) for i,(a,x) in enumerate(ins) if re.search(r"\bCALL",x)]
    tgt=None
    for i,x in calls:
        m=re.search(r"0x([0-9a-f]+)",x)
        if m: tgt=int(m.group(1),16)
    split=len(ins)
    if tgt is not None:
        for i,(a,x) in enumerate(ins):
            if a==tgt: split=i
    body=ins[:split]
    stl=[i for i,(a,x) in enumerate(body) if re.search(r"\bSTL",x)]; ldl=[i for i,(a,x) in enumerate(body) if re.search(r"\bLDL",x)]
    print("%-40s kernel body n=%d (callee %d) STL=%d LDL=%d calls=%s"%(f[:40],len(body),len(ins)-split,len(stl),len(ldl),[i for i,_ in calls]))
    if calls:
        c=[i for i,_ in calls]
        near=lambda i: min(abs(i-k) for k in c)
        print("    STL within 60 of the CALL: %d, elsewhere: %d ; LDL near: %d, elsewhere: %d"%(sum(near(i)<=60 for i in stl),sum(near(i)>60 for i in stl),sum(near(i)<=60 for i in ldl),sum(near(i)>60 for i in ldl)))
