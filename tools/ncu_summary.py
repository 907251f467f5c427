#!/usr/bin/env python
"""Summarise an .ncu-rep (one kernel capture, `ncu --set full --import-source on`) into a
small text file for profiles/: headline metrics + instructions per source line.

    python tools/ncu_summary.py gpurun_out/prof_fused.ncu-rep ENV_STEPS_PER_LAUNCH > profiles/<name>.txt
"""
import csv
import io
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "launch__waves_per_multiprocessor",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__inst_executed.sum.per_cycle_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "smsp__sass_average_branch_targets_threads_uniform.pct", "smsp__sass_branch_targets_threads_divergent.sum",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_bytes.sum", "l1tex__t_bytes.sum", "sm__cycles_elapsed.max"]


def ncu(rep, *args):
    return subprocess.run(["ncu", "-i", rep, *args], capture_output=True, text=True).stdout


def main():
    rep = sys.argv[1]
    env_steps = float(sys.argv[2]) if len(sys.argv) > 2 else None
    rows = list(csv.reader(io.StringIO(ncu(rep, "--page", "raw", "--csv"))))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        d = dict(zip(hdr, r))
        print("kernel:", d.get("Kernel Name"))
        for k in KEYS:
            if k in d:
                print("  %-62s %s %s" % (k, d[k], units[hdr.index(k)]))
        if env_steps:
            inst = float(d["smsp__inst_executed.sum"])
            print("  warp instructions per env-step: %.1f" % (inst / env_steps))
            dram = 0.0
            for k in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                u = units[hdr.index(k)].lower()
                dram += float(d[k]) * {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}[u]
            print("  DRAM bytes per env-step: %.1f (per launch %.0f)" % (dram / env_steps, dram))
    rows = list(csv.reader(io.StringIO(ncu(rep, "--page", "source", "--csv", "--print-source", "cuda,sass"))))
    agg, cur = [], None
    for r in rows:
        if len(r) >= 2 and r[0] == "File Path":
            cur = r[1].split("/")[-1]
        elif len(r) > 8 and r[0].isdigit():
            try:
                agg.append((int(r[7]), int(r[6]), cur, int(r[0]), r[1].strip()[:100]))
            except ValueError:
                pass
    tot = sum(a[0] for a in agg) or 1
    samp = sum(a[1] for a in agg) or 1
    print("\ninstructions executed per source line (inlined callees attributed to their own lines), top 40")
    print("  %8s %6s %7s  %s" % ("inst/step" if env_steps else "inst", "inst%", "stall%", "line"))
    for a in sorted(agg, reverse=True)[:40]:
        per = a[0] / env_steps if env_steps else a[0]
        print("  %8.1f %5.1f%% %6.1f%%  %s:%d  %s" % (per, 100.0 * a[0] / tot, 100.0 * a[1] / samp, a[2], a[3], a[4]))


if __name__ == "__main__":
    main()
