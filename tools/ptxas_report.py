#!/usr/bin/env python3
"""Summarise ptxas -v logs: kernel, registers, stack, spills.  usage: ptxas_report.py LOG... [--grep S]"""
import re, subprocess, sys
files = [a for a in sys.argv[1:] if not a.startswith("--")]
pat = None
if "--grep" in sys.argv:
    pat = sys.argv[sys.argv.index("--grep") + 1]
    files = [f for f in files if f != pat]
for f in files:
    name = None
    stack = ""
    for line in open(f):
        m = re.search(r"Compiling entry function '([^']+)'", line)
        if m:
            name = m.group(1)
        m = re.search(r"(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads", line)
        if m:
            stack = "stack %s spill %s/%s" % m.groups()
        m = re.search(r"Used (\d+) registers", line)
        if m and name:
            dem = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip()
            dem = re.sub(r"\(.*", "", dem)
            if pat is None or pat in dem:
                print("%-48s regs %3s  %s" % (dem[-48:], m.group(1), stack))
            name = None
