#!/usr/bin/env python3
"""Build oracle/_ref/libaomref.so from the UNMODIFIED reference sources where they lie.

TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product path.

Recipe (no cmake, no copy of reference sources into this repo):
  1. config/aom_config.h  <- the default values listed in the reference's
     build/cmake/aom_config_defaults.cmake, target "generic" (pure C, no SIMD).
  2. config/{av1,aom_dsp,aom_scale}_rtcd.h <- the reference's own build/cmake/rtcd.pl
     run on its *_rtcd_defs.pl with --arch=generic (every symbol becomes a macro for
     its _c flavour; this is what `-DAOM_TARGET_CPU=generic` produces).
  3. gcc -O2 -shared on the handful of leaf sources of the hot path (list below),
     compiled in place from /root/reference.  Encoder-control functions that those
     files reference but that the hot path never calls (compound/OBMC/warp/hash
     helpers) are listed with `nm` after a first link and given abort() stubs in
     _ref/auto_stubs.c so that the library loads with RTLD_NOW.

Outputs go ONLY to oracle/_ref/ (git-ignored, travels to the GPU box as a built .so).
If /root/reference is absent (GPU box) this script is a no-op.
"""
import os
import re
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("AOM_REFERENCE", "/root/reference")
OUT = os.path.join(HERE, "_ref")

# Leaf sources of the hot path compiled where they lie.
SOURCES = [
    "av1/encoder/av1_fwd_txfm1d.c",
    "av1/encoder/av1_fwd_txfm2d.c",
    "av1/common/av1_txfm.c",
    "av1/encoder/hybrid_fwd_txfm.c",
    "av1/encoder/av1_quantize.c",
    "aom_dsp/quantize.c",
    "av1/common/scan.c",
    "av1/common/quant_common.c",
    "av1/common/common_data.c",
    "aom_dsp/sad.c",
    "aom_dsp/variance.c",
    "aom_dsp/subtract.c",
    "aom_dsp/sse.c",
    "aom_dsp/avg.c",
    "aom_dsp/aom_convolve.c",
    "av1/common/convolve.c",
    "av1/encoder/mcomp.c",
    "av1/encoder/reconinter_enc.c",
    "av1/encoder/rdopt.c",
]

OVERRIDES = {
    "INLINE": "inline",
    "HAVE_PTHREAD_H": "1",
    "HAVE_UNISTD_H": "1",
    "CONFIG_RUNTIME_CPU_DETECT": "0",
}


def write_config(cfgdir):
    text = open(os.path.join(REF, "build/cmake/aom_config_defaults.cmake")).read()
    vals = {}
    for m in re.finditer(r'set_aom_(?:detect|config)_var\(\s*(\w+)\s+("[^"]*"|\S+)', text):
        vals[m.group(1)] = m.group(2).strip('"')
    vals.update(OVERRIDES)
    keep = ("ARCH_", "HAVE_", "CONFIG_", "INLINE", "DECODE_", "FORCE_", "STATIC_")
    with open(os.path.join(cfgdir, "aom_config.h"), "w") as f:
        f.write("/* written by oracle/mk_ref.py from aom_config_defaults.cmake (generic target) */\n")
        f.write("#ifndef AOM_CONFIG_H_\n#define AOM_CONFIG_H_\n")
        for k, v in vals.items():
            if k.startswith(keep):
                f.write("#define %s %s\n" % (k, v))
        f.write("#endif\n")


def write_rtcd(cfgdir):
    for sym, defs in (
        ("av1_rtcd", "av1/common/av1_rtcd_defs.pl"),
        ("aom_dsp_rtcd", "aom_dsp/aom_dsp_rtcd_defs.pl"),
        ("aom_scale_rtcd", "aom_scale/aom_scale_rtcd.pl"),
    ):
        out = subprocess.run(
            ["perl", os.path.join(REF, "build/cmake/rtcd.pl"), "--arch=generic",
             "--sym=" + sym, "--config=" + os.path.join(cfgdir, "aom_config.h"),
             os.path.join(REF, defs)],
            check=True, capture_output=True, text=True).stdout
        with open(os.path.join(cfgdir, sym + ".h"), "w") as f:
            f.write(out)


def main():
    if not os.path.isdir(REF):
        print("mk_ref: %s absent; keeping any prebuilt oracle/_ref" % REF)
        return 0
    cfgdir = os.path.join(OUT, "config")
    os.makedirs(cfgdir, exist_ok=True)
    lib = os.path.join(OUT, "libaomref.so")
    stubs = os.path.join(OUT, "auto_stubs.c")
    newest = max(os.path.getmtime(__file__), os.path.getmtime(os.path.join(HERE, "ref_shim.c")),
                 os.path.getmtime(os.path.join(HERE, "av1o_me.h")))
    if os.path.exists(lib) and os.path.getmtime(lib) > newest and "--force" not in sys.argv:
        print("mk_ref: up to date")
        return 0
    write_config(cfgdir)
    write_rtcd(cfgdir)
    cmd = ["gcc", "-O2", "-fPIC", "-shared", "-std=c99", "-D_GNU_SOURCE", "-DNDEBUG", "-w",
           "-I" + REF, "-I" + OUT]
    objs = []
    objdir = os.path.join(OUT, "obj")
    os.makedirs(objdir, exist_ok=True)
    procs = []
    srcs = [os.path.join(REF, s) for s in SOURCES] + [os.path.join(HERE, "ref_shim.c")]
    for s in srcs:
        o = os.path.join(objdir, os.path.relpath(s, "/").replace("/", "_")[:-2] + ".o")
        objs.append(o)
        procs.append(subprocess.Popen(cmd[:1] + ["-c"] + [c for c in cmd[1:] if c != "-shared"]
                                      + ["-I" + HERE, s, "-o", o]))
    for p in procs:
        if p.wait() != 0:
            raise SystemExit("mk_ref: compile failed")
    link = ["gcc", "-shared", "-o", lib] + objs
    open(stubs, "w").write("/* empty */\n")
    subprocess.run(link + [stubs, "-fPIC", "-lm", "-lpthread"], check=True)
    undef = []
    for line in subprocess.run(["nm", "-D", lib], check=True, capture_output=True,
                               text=True).stdout.splitlines():
        parts = line.split()
        if len(parts) == 2 and parts[0] == "U" and "@" not in parts[1] and not parts[1].startswith("_"):
            undef.append(parts[1])
    with open(stubs, "w") as f:
        f.write("/* abort() stubs for control-plane symbols the hot path never calls */\n")
        f.write("#include <stdio.h>\n#include <stdlib.h>\n")
        for u in undef:
            f.write('void %s(void) { fprintf(stderr, "aomref stub called: %s\\n"); abort(); }\n' % (u, u))
    subprocess.run(link + [stubs, "-fPIC", "-lm", "-lpthread"], check=True)
    print("mk_ref: built", lib)
    return 0


if __name__ == "__main__":
    sys.exit(main())
