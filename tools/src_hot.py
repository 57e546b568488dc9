#!/usr/bin/env python3
"""Summarise `ncu --page source --csv --print-source cuda,sass`: per kernel, the CUDA source
lines that execute the most warp instructions / collect the most stall samples."""
import collections
import csv
import sys


def main():
    path, topn = sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 30
    want = sys.argv[3] if len(sys.argv) > 3 else ""
    fn = fpath = None
    hdr = None
    agg = collections.defaultdict(lambda: collections.defaultdict(lambda: [0.0, 0.0, ""]))
    with open(path, newline="") as f:
        for row in csv.reader(f):
            if not row:
                continue
            if row[0] == "File Path":
                fpath = row[1].split("/")[-1]
            elif row[0] == "Function Name":
                fn = row[1]
            elif row[0] == "Line No":
                hdr = {h: i for i, h in enumerate(row)}
                i_inst = row.index("Instructions Executed")
                i_samp = row.index("# Samples")
            elif hdr and row[0].isdigit():
                try:
                    a = agg[fn][(fpath, int(row[0]))]
                    a[0] += float(row[i_inst] or 0)
                    a[1] += float(row[i_samp] or 0)
                    a[2] = row[1].strip()[:90]
                except ValueError:
                    pass
    for fn, lines in agg.items():
        if want not in fn:
            continue
        ti = sum(v[0] for v in lines.values())
        ts = sum(v[1] for v in lines.values())
        print("=" * 110)
        print(fn, " warp-instr", int(ti), " samples", int(ts))
        print("-- by instructions executed")
        for (fp, ln), v in sorted(lines.items(), key=lambda kv: -kv[1][0])[:topn]:
            print("  %5.2f%% inst %5.2f%% samp  %s:%d  %s" % (100 * v[0] / max(ti, 1), 100 * v[1] / max(ts, 1), fp, ln, v[2]))
        print("-- by stall samples")
        for (fp, ln), v in sorted(lines.items(), key=lambda kv: -kv[1][1])[:topn]:
            print("  %5.2f%% samp %5.2f%% inst  %s:%d  %s" % (100 * v[1] / max(ts, 1), 100 * v[0] / max(ti, 1), fp, ln, v[2]))


if __name__ == "__main__":
    main()
