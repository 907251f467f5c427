#!/usr/bin/env python
"""profiles/traffic.json from the round's ncu captures (`ncu --set full`): per kernel the DRAM bytes and warp
instructions per unit of work, issue-slot and ALU-pipe utilisation — what bench.py reports next to the live timings.
    python tools/make_traffic.py TAG      (reads gpurun_out/TAG_{fused,fused_M,reset,step_sampled}.ncu-rep)"""
import csv, io, json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1]
UNIT = {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}


def raw(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    h, u, d = rows[0], rows[1], rows[2]
    return {k: (v, uu) for k, uu, v in zip(h, u, d)}


def entry(name, units, unit_name, extra=None):
    rep = os.path.join(ROOT, "gpurun_out", "%s_%s.ncu-rep" % (tag, name))
    if not os.path.exists(rep):
        return None
    m = raw(rep)
    dram = sum(float(m[k][0]) * UNIT[m[k][1].lower()] for k in ("dram__bytes_read.sum", "dram__bytes_write.sum"))
    inst = float(m["smsp__inst_executed.sum"][0])
    e = {"kernel": m["Kernel Name"][0], "duration_us_under_ncu": float(m["gpu__time_duration.sum"][0]) *
         {"us": 1, "ms": 1e3, "ns": 1e-3, "s": 1e6}.get(m["gpu__time_duration.sum"][1].lower().replace("second", "s").replace("usecond", "us"), 1),
         "dram_bytes_per_launch": dram, "dram_bytes_per_%s" % unit_name: dram / units,
         "warp_inst_per_%s" % unit_name: inst / units,
         "issue_active_frac": float(m["smsp__issue_active.avg.pct_of_peak_sustained_active"][0]) / 100,
         "alu_pipe_frac": float(m["sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"][0]) / 100,
         "thread_efficiency": float(m["smsp__thread_inst_executed_per_inst_executed.ratio"][0]) / 32,
         "registers": int(float(m["launch__registers_per_thread"][0])),
         "source": "profiles/%s_%s.txt (ncu --set full --clock-control none)" % (tag, name)}
    e.update(extra or {})
    return e


n = 65536
out = {"k_step_fused": entry("fused", n * 1024, "env_step", {"inner": 1024, "envs": n}),
       "k_step_fused_M": entry("fused_M", n * 256, "env_step", {"inner": 256, "envs": n}),
       "k_reset": entry("reset", n, "reset", {"envs": n}),
       "k_step_single": entry("step_sampled", n, "env_step", {"envs": n})}
out = {k: v for k, v in out.items() if v}
json.dump(out, open(os.path.join(ROOT, "profiles", "traffic.json"), "w"), indent=1)
print(json.dumps(out, indent=1))
