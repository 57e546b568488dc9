#!/usr/bin/env python3
"""Summarise `ncu --page source --csv --print-source sass` exports: per kernel, executed
instruction totals, opcode histogram weighted by executions, top stall sites."""
import csv, sys, collections, re

def kernels(path):
    cur, hdr, rows = None, None, []
    with open(path, newline="") as f:
        for row in csv.reader(f):
            if not row: continue
            if row[0] == "Kernel Name":
                if cur: yield cur, hdr, rows
                cur, hdr, rows = row[1], None, []
            elif row[0] == "Address":
                hdr = row
            elif hdr:
                rows.append(row)
    if cur: yield cur, hdr, rows

def main():
    path = sys.argv[1]
    topn = int(sys.argv[2]) if len(sys.argv) > 2 else 25
    for name, hdr, rows in kernels(path):
        ix = {h: i for i, h in enumerate(hdr)}
        def col(r, h):
            try: return float(r[ix[h]] or 0)
            except Exception: return 0.0
        tot_exec = sum(col(r, "Instructions Executed") for r in rows)
        tot_samp = sum(col(r, "# Samples") for r in rows)
        print("=" * 100)
        print(name, " SASS lines", len(rows), " warp-instr executed", int(tot_exec), " samples", int(tot_samp))
        ops = collections.Counter(); samp = collections.Counter()
        for r in rows:
            m = re.match(r"\s*(@!?U?P\d+\s+)?([A-Z0-9_.]+)", r[ix["Source"]])
            op = m.group(2) if m else "?"
            op = ".".join(op.split(".")[:2])
            ops[op] += col(r, "Instructions Executed"); samp[op] += col(r, "# Samples")
        print("-- opcode: executed share / sample share")
        for op, n in ops.most_common(topn):
            print(f"  {op:24s} {n/tot_exec*100:6.2f}%  {samp[op]/max(tot_samp,1)*100:6.2f}%")
        print("-- top stall sites (samples)")
        order = sorted(range(len(rows)), key=lambda i: -col(rows[i], "# Samples"))[:topn]
        for i in order:
            r = rows[i]
            print(f"  {int(col(r,'# Samples')):7d} {col(r,'# Samples')/max(tot_samp,1)*100:5.2f}%  exec {int(col(r,'Instructions Executed')):9d}  {r[ix['Source']][:90]}")

if __name__ == "__main__":
    main()
