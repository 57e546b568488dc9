#!/usr/bin/env python3
"""Summarise an ncu --set full report (exported with `ncu -i X.ncu-rep --page raw --csv`)
into a markdown table for profiles/.  usage: ncu_summary.py raw.csv frames_per_launch title"""
import csv
import re
import sys

rows = list(csv.reader(open(sys.argv[1])))
F = int(sys.argv[2])
title = sys.argv[3] if len(sys.argv) > 3 else "ncu summary"
hdr, units, data = rows[0], rows[1], rows[2:]
idx = {h: i for i, h in enumerate(hdr)}
# ncu picks a unit per column (byte/Kbyte/Mbyte/Gbyte, ns/us/ms/s): normalise to MB and us
SCALE = {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3, "Tbyte": 1e6,
         "ns": 1e-3, "nsecond": 1e-3, "us": 1.0, "usecond": 1.0, "ms": 1e3, "msecond": 1e3,
         "s": 1e6, "second": 1e6}
print("# %s\n" % title)
print("| kernel | us | dram rd MB | dram wr MB | alg MB | dram % | alu % | fma % | issue % | regs | "
      "warps active % | thread-inst/sample | no_instr | long_sb | math_throttle |")
print("|---|---|---|---|---|---|---|---|---|---|---|---|---|---|---|")
for r in data:
    name = r[idx["Kernel Name"]]
    g = lambda k: float(r[idx[k]].replace(",", "")) * SCALE.get(units[idx[k]], 1.0)
    st = lambda k: g("smsp__average_warps_issue_stalled_%s_per_issue_active.ratio" % k)
    m = re.search(r"txq_(?:small_)?kernel<(\d+), (\d+), (\d+)", name)
    if m:
        W, H = int(m.group(1)), int(m.group(2))
        nb = sum((w // W) * (h // H) for (w, h) in [(1920, 1080), (960, 540), (960, 540)]) * F
        alg = nb * (W * H * 2 + min(W, 32) * min(H, 32) * 12 + 2) / 1e6
        samples = nb * W * H
        label = "txq%s %dx%d" % ("_small" if "small" in name else "", W, H)
    else:
        alg, samples, label = float("nan"), float("nan"), name.split("(")[0][-40:]
    print("| %s | %.1f | %.1f | %.1f | %.1f | %.1f | %.1f | %.1f | %.1f | %d | %.1f | %.1f | %.2f | %.2f | %.2f |" % (
        label, g("gpu__time_duration.sum"), g("dram__bytes_read.sum"), g("dram__bytes_write.sum"), alg,
        g("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
        g("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"),
        g("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active"),
        g("smsp__issue_active.avg.pct_of_peak_sustained_active"), int(g("launch__registers_per_thread")),
        g("sm__warps_active.avg.pct_of_peak_sustained_active"),
        g("smsp__inst_executed.sum") * 32 / samples if samples == samples else float("nan"),
        st("no_instruction"), st("long_scoreboard"), st("math_pipe_throttle")))
