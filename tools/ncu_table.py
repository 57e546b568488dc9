#!/usr/bin/env python3
"""Generic summary of an `ncu --page raw --csv` export: one row per kernel launch with the counters the
round's analysis uses.  usage: ncu_table.py raw.csv [--dedupe]  (--dedupe keeps the last launch per kernel name)"""
import csv
import re
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr, units, data = rows[0], rows[1], rows[2:]
idx = {h: i for i, h in enumerate(hdr)}
SCALE = {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3, "Tbyte": 1e6,
         "ns": 1e-3, "nsecond": 1e-3, "us": 1.0, "usecond": 1.0, "ms": 1e3, "msecond": 1e3, "s": 1e6, "second": 1e6}


def g(r, k):
    if k not in idx or r[idx[k]] in ("", "n/a"):
        return float("nan")
    return float(r[idx[k]].replace(",", "")) * SCALE.get(units[idx[k]], 1.0)


if "--dedupe" in sys.argv:
    last = {}
    for r in data:
        last[r[idx["Kernel Name"]]] = r
    data = list(last.values())
print("| kernel | us | dram rd MB | dram wr MB | dram % | L1 ld sectors/req | L2->L1 MB | alu % | fma % | issue % | regs | "
      "smem KB/CTA | warps active % | long_sb | short_sb | barrier | math_thr | no_instr | wait |")
print("|" + "---|" * 19)
for r in data:
    name = re.sub(r"^void ", "", r[idx["Kernel Name"]]).split("(")[0]
    st = lambda k: g(r, "smsp__average_warps_issue_stalled_%s_per_issue_active.ratio" % k)
    req = g(r, "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum")
    sec = g(r, "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum")
    print("| %s | %.1f | %.1f | %.1f | %.1f | %.2f | %.1f | %.1f | %.1f | %.1f | %d | %.1f | %.1f | %.2f | %.2f | %.2f | %.2f | %.2f | %.2f |" % (
        name[-44:], g(r, "gpu__time_duration.sum"), g(r, "dram__bytes_read.sum"), g(r, "dram__bytes_write.sum"),
        g(r, "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
        sec / req if req == req and req > 0 else float("nan"),
        g(r, "lts__t_sectors_srcunit_tex.sum") * 32 / 1e6,
        g(r, "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"),
        g(r, "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active"),
        g(r, "smsp__issue_active.avg.pct_of_peak_sustained_active"), int(g(r, "launch__registers_per_thread")),
        (g(r, "launch__shared_mem_per_block_dynamic") + g(r, "launch__shared_mem_per_block_static")) * 1e3
        if True else 0,
        g(r, "sm__warps_active.avg.pct_of_peak_sustained_active"),
        st("long_scoreboard"), st("short_scoreboard"), st("barrier"), st("math_pipe_throttle"), st("no_instruction"),
        st("wait")))
