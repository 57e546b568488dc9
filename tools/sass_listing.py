#!/usr/bin/env python3
"""profiles/sass/: trimmed SASS listings of one representative kernel per class from the shipped
libb200av1.so (cuobjdump -sass), plus an opcode census of the whole library (TMA / bulk-copy / packed
integer instructions that the design claims).  usage: sass_listing.py LIB OUTDIR"""
import collections
import os
import re
import subprocess
import sys

lib, out = sys.argv[1], sys.argv[2]
os.makedirs(out, exist_ok=True)
txt = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
funcs, cur = {}, None
for line in txt.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        cur = m.group(1)
        funcs[cur] = []
    elif cur and re.match(r"\s+/\*[0-9a-f]{4,}\*/", line):
        funcs[cur].append(re.sub(r"\s*/\* 0x[0-9a-f]+ \*/\s*$", "", line).rstrip())
dem = {f: subprocess.run(["c++filt", f], capture_output=True, text=True).stdout.strip().split("(")[0] for f in funcs}
want = {"txq_kernel_32x32_residual": "txq_kernel<32, 32, 0>", "txq_kernel_16x16_pixels16": "txq_kernel<16, 16, 2>",
        "txq_small_kernel_4x4_residual": "txq_small_kernel<4, 4, 0>", "itx_kernel_16x16_u8": "itx_kernel<16, 16, 1>",
        "itx_kernel_32x32_u16": "itx_kernel<32, 32, 2>", "subpel_kernel": "subpel_kernel",
        "hadamard_kernel_16": "hadamard_kernel<16, 0>", "compound_kernel": "compound_kernel",
        "rate_est_kernel": "rate_est_kernel", "pix_kernel": "pix_kernel"}
for name, pat in want.items():
    for f, d in dem.items():
        if d.endswith(pat):
            with open(os.path.join(out, name + ".sass"), "w") as fh:
                fh.write("// %s  (%d SASS instructions)\n" % (d, len(funcs[f])))
                fh.write("\n".join(funcs[f]) + "\n")
            break
# the full-pel kernel is ~45k instructions: keep the lines that carry the claims (TMA, mbarrier, the
# packed SAD instructions) with a little context, and the census
census = collections.OrderedDict()
claims = re.compile(r"UTMALDG|UTMASTG|UBLKCP|SYNCS|VABSDIFF4|VIMNMX\.U16x2|IDP\.2A|IDP\.4A|LDGSTS|REDUX|ATOMS|LDSM|UTMAPF")
with open(os.path.join(out, "census.txt"), "w") as fh:
    fh.write("opcode census of %s (cuobjdump -sass): instructions per kernel that carry the design's claims\n\n" % os.path.basename(lib))
    tot = collections.Counter()
    for f, lines in funcs.items():
        c = collections.Counter()
        for l in lines:
            m = claims.search(l)
            if m:
                c[m.group(0)] += 1
        tot.update(c)
        if c:
            fh.write("%-60s %6d instr  %s\n" % (dem[f][-60:], len(lines), "  ".join("%s=%d" % kv for kv in sorted(c.items()))))
    fh.write("\nwhole library: %s\n" % "  ".join("%s=%d" % kv for kv in sorted(tot.items())))
for f, d in dem.items():
    if d.endswith("fullpel_kernel<0>"):
        lines = funcs[f]
        keep = set()
        for i, l in enumerate(lines):
            if re.search(r"UTMALDG|SYNCS|VIMNMX\.U16x2|IDP\.2A", l):
                keep.update(range(max(0, i - 3), min(len(lines), i + 4)))
        with open(os.path.join(out, "fullpel_kernel_nstep_mesh_excerpt.sass"), "w") as fh:
            fh.write("// %s: %d SASS instructions; excerpt = TMA / mbarrier / packed 16-bit SAD sites +-3 lines\n" % (d, len(lines)))
            last = -2
            for i in sorted(keep):
                if i != last + 1:
                    fh.write("        ...\n")
                fh.write(lines[i] + "\n")
                last = i
