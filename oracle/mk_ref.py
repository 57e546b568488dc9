#!/usr/bin/env python3
"""Build the UNMODIFIED reference (libaom, aom-av1-lavish fork) from its sources where they lie.

TEST / BASELINE INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product path.

Outputs (only under oracle/_ref/, git-ignored, they travel to the GPU box as built files):
  libaomref.so        generic C target (-DAOM_TARGET_CPU=generic equivalent): the whole library
                      (common + encoder + decoder) + oracle/ref_shim.c.  This is the parity oracle:
                      every `_c` function of the hot path, av1_full_pixel_search, the sub-pel
                      trees, the temporal-filter / TPL helpers.
  libaomref_avx2.so   same sources + the x86 INTRINSICS files (sse2 .. avx2, gcc -m flags), rtcd
                      generated with --arch=x86_64 and run-time CPU detection.  No assembler is
                      available in this image, so the kernels that exist only as .asm (sad_sse2.asm,
                      subpel_variance_sse2.asm, ...) are aliased to their `_c` flavour at link time
                      (ld --defsym) -- the "best-effort SIMD baseline B'" of SURVEY.md section 8d.
  aomenc, aomenc_avx2 the reference encoder CLI on top of each (BASELINE configs[0]).

Recipe (no cmake, no copy of reference sources into this repo):
  1. config/aom_config.h  <- the default values of build/cmake/aom_config_defaults.cmake with the
     per-flavour target flags; config/aom_version.h, aom_config.c written here.
  2. config/{av1,aom_dsp,aom_scale}_rtcd.h <- the reference's own build/cmake/rtcd.pl on its
     *_rtcd_defs.pl (--arch=generic | x86_64).
  3. gcc -O2 -c on every .c of aom/, aom_dsp/, aom_mem/, aom_ports/, aom_scale/, aom_util/, av1/
     (arch sub-directories only for the avx2 flavour), in place from /root/reference.

If /root/reference is absent (GPU box) this script is a no-op and the prebuilt files are used.
"""
import concurrent.futures
import glob
import os
import re
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("AOM_REFERENCE", "/root/reference")
OUT = os.path.join(HERE, "_ref")

LIB_DIRS = ["aom/src", "aom_dsp", "aom_mem", "aom_ports", "aom_scale", "aom_scale/generic", "aom_util",
            "av1", "av1/common", "av1/encoder", "av1/decoder", "aom_dsp/flow_estimation", "third_party/vector",
            "third_party/fastfeat"]
LIB_EXTRA = ["common/args_helper.c"]
# sources that belong to configurations this build leaves off, or to other CPU families
EXCLUDE = {"aom_ports/arm_cpudetect.c", "aom_ports/ppc_cpudetect.c", "av1/encoder/tune_vmaf.c",
           "av1/encoder/tune_butteraugli.c", "av1/encoder/blockiness.c", "av1/encoder/saliency_map.c",
           "av1/encoder/sorting_network.h", "aom_dsp/psnrhvs.c", "aom_dsp/ssim.c", "aom_dsp/fastssim.c",
           "aom_dsp/butteraugli.c", "aom_dsp/vmaf.c", "av1/decoder/accounting.c", "av1/decoder/inspection.c",
           "aom_util/debug_util.c", "av1/encoder/av1_temporal_denoiser.c", "aom_dsp/noise_model.c",
           "aom_dsp/noise_util.c"}
APP_SOURCES = ["apps/aomenc.c", "common/args.c", "common/ivfenc.c", "common/md5_utils.c", "examples/encoder_util.c", "common/rawenc.c",
               "common/tools_common.c", "common/warnings.c", "common/y4minput.c", "common/y4menc.c",
               "common/ivfdec.c", "common/obudec.c", "common/av1_config.c", "stats/aomstats.c",
               "stats/rate_hist.c", "av1/arg_defs.c"]
X86_EXCLUDE = {"av1/encoder/x86/av1_temporal_denoiser_sse2.c"}  # CONFIG_AV1_TEMPORAL_DENOISING=0
X86_DIRS = ["aom_dsp/x86", "av1/common/x86", "av1/encoder/x86", "aom_dsp/flow_estimation/x86"]

COMMON_OVERRIDES = {
    "INLINE": "inline", "HAVE_PTHREAD_H": "1", "HAVE_UNISTD_H": "1", "CONFIG_WEBM_IO": "0",
    "CONFIG_LIBYUV": "0", "CONFIG_TUNE_VMAF": "0", "CONFIG_TUNE_BUTTERAUGLI": "0",
    "CONFIG_TENSORFLOW_LITE": "0", "CONFIG_INTERNAL_STATS": "0", "CONFIG_DENOISE": "0",
    "CONFIG_AV1_TEMPORAL_DENOISING": "0", "CONFIG_THREE_PASS": "0", "CONFIG_SALIENCY_MAP": "0",
}
FLAVOURS = {
    "generic": {"arch": "generic", "cfg": {"CONFIG_RUNTIME_CPU_DETECT": "0", "AOM_ARCH_X86_64": "0",
                                           "ARCH_X86_64": "0"}, "suffix": ""},
    "avx2": {"arch": "x86_64",
             "cfg": {"CONFIG_RUNTIME_CPU_DETECT": "1", "AOM_ARCH_X86_64": "1", "ARCH_X86_64": "1",
                     "HAVE_MMX": "1", "HAVE_SSE": "1", "HAVE_SSE2": "1", "HAVE_SSE3": "1", "HAVE_SSSE3": "1",
                     "HAVE_SSE4_1": "1", "HAVE_SSE4_2": "1", "HAVE_AVX": "1", "HAVE_AVX2": "1"},
             "suffix": "_avx2"},
}
ISA_FLAGS = [("avx2", "-mavx2"), ("avx", "-mavx"), ("sse4_2", "-msse4.2"), ("sse4", "-msse4.1"), ("sse42", "-msse4.2"),
             ("ssse3", "-mssse3"), ("sse3", "-msse3"), ("sse2", "-msse2"), ("sse", "-msse"), ("mmx", "-mmmx")]


def write_if_changed(path, text):
    """Generated headers keep their mtime when nothing changed, so the object cache stays valid."""
    if os.path.exists(path) and open(path).read() == text:
        return
    with open(path, "w") as f:
        f.write(text)


def write_config(cfgdir, flavour):
    text = open(os.path.join(REF, "build/cmake/aom_config_defaults.cmake")).read()
    vals = {}
    for m in re.finditer(r'set_aom_(?:detect|config)_var\(\s*(\w+)\s+("[^"]*"|\S+)', text):
        vals[m.group(1)] = m.group(2).strip('"')
    vals.update(COMMON_OVERRIDES)
    vals.update(FLAVOURS[flavour]["cfg"])
    keep = ("ARCH_", "AOM_ARCH_", "HAVE_", "CONFIG_", "INLINE", "DECODE_", "FORCE_", "STATIC_")
    text = "/* written by oracle/mk_ref.py from aom_config_defaults.cmake (%s target) */\n" % flavour
    text += "#ifndef AOM_CONFIG_H_\n#define AOM_CONFIG_H_\n"
    for k, v in vals.items():
        if k.startswith(keep):
            text += "#define %s %s\n" % (k, v)
    write_if_changed(os.path.join(cfgdir, "aom_config.h"), text + "#endif\n")
    write_if_changed(os.path.join(cfgdir, "aom_version.h"), "#ifndef AOM_VERSION_H_\n#define AOM_VERSION_H_\n#define VERSION_MAJOR 3\n#define VERSION_MINOR 6\n"
                "#define VERSION_PATCH 0\n#define VERSION_EXTRA \"lavish-ref\"\n"
                "#define VERSION_PACKED ((VERSION_MAJOR << 16) | (VERSION_MINOR << 8) | (VERSION_PATCH))\n"
                "#define VERSION_STRING_NOSP \"3.6.0-lavish-ref\"\n#define VERSION_STRING \" 3.6.0-lavish-ref\"\n#endif\n")
    write_if_changed(os.path.join(cfgdir, "aom_config.c"), '#include "aom/aom_codec.h"\nstatic const char *const cfg = "oracle/mk_ref.py %s";\n'
                "const char *aom_codec_build_config(void) { return cfg; }\n" % flavour)


def write_rtcd(cfgdir, arch):
    for sym, defs in (("av1_rtcd", "av1/common/av1_rtcd_defs.pl"), ("aom_dsp_rtcd", "aom_dsp/aom_dsp_rtcd_defs.pl"),
                      ("aom_scale_rtcd", "aom_scale/aom_scale_rtcd.pl")):
        out = subprocess.run(["perl", os.path.join(REF, "build/cmake/rtcd.pl"), "--arch=" + arch, "--sym=" + sym,
                              "--config=" + os.path.join(cfgdir, "aom_config.h"), os.path.join(REF, defs)],
                             check=True, capture_output=True, text=True).stdout
        write_if_changed(os.path.join(cfgdir, sym + ".h.gen"), out)


def isa_flag(path):
    base = os.path.basename(path)
    for tag, flag in ISA_FLAGS:
        if re.search(r"(^|_)%s(_|\.)" % tag, base):
            return [flag]
    return ["-msse2"]


def compile_all(srcs, objdir, incs, defs, hdrdir):
    os.makedirs(objdir, exist_ok=True)
    jobs = []
    for s, extra in srcs:
        o = os.path.join(objdir, os.path.relpath(s, "/").replace("/", "_")[:-2] + ".o")
        jobs.append((s, o, extra))

    newest_hdr = max([os.path.getmtime(__file__)] + [os.path.getmtime(p) for p in glob.glob(os.path.join(hdrdir, "*.h"))]
                     + [os.path.getmtime(p) for p in glob.glob(os.path.join(HERE, "*.h"))])

    def one(job):
        s, o, extra = job
        if os.path.exists(o) and os.path.getmtime(o) > max(os.path.getmtime(s), newest_hdr):
            return o, 0, ""
        r = subprocess.run(["gcc", "-O2", "-fPIC", "-ffunction-sections", "-std=gnu99", "-D_GNU_SOURCE", "-DNDEBUG", "-w", "-c"] + defs +
                           extra + incs + [s, "-o", o], capture_output=True, text=True)
        return o, r.returncode, r.stderr
    objs, failed = [], []
    with concurrent.futures.ThreadPoolExecutor(max_workers=os.cpu_count() or 4) as ex:
        for (o, rc, err), job in zip(ex.map(one, jobs), jobs):
            if rc != 0:
                failed.append((job[0], err))
            else:
                objs.append(o)
    return objs, failed


EXT_RE = r"_(mmx|sse|sse2|sse3|ssse3|sse4_1|sse4_2|avx|avx2)$"


def asm_dependent(objs):
    """Functions of the intrinsics objects that reach a helper which exists only as assembler
    (aom_sub_pixel_variance8xh_sse2, aom_filter_block1d8_h8_ssse3, aom_highbd_calc16x16var_sse2 ...):
    relocations per function section (-ffunction-sections), transitive closure."""
    refs, defined, wanted = {}, set(), set()
    for o in objs:
        for line in subprocess.run(["nm", o], capture_output=True, text=True).stdout.splitlines():
            parts = line.split()
            if len(parts) == 3 and parts[1] in "Tt":
                defined.add(parts[2])
            elif len(parts) == 2 and parts[0] == "U":
                wanted.add(parts[1])
        cur = None
        for line in subprocess.run(["objdump", "-dr", "--no-show-raw-insn", o], capture_output=True,
                                   text=True).stdout.splitlines():
            m = re.match(r"^[0-9a-f]+ <([\w.]+)>:", line)
            if m:
                cur = m.group(1)
                refs.setdefault(cur, set())
                continue
            m = re.search(r"R_X86_64_\w+\s+([\w.]+)", line)
            if m and cur:
                t = m.group(1)
                refs[cur].add(t[len(".text."):] if t.startswith(".text.") else t)
    asm_only = {w for w in wanted - defined if re.search(EXT_RE, w)}
    tainted = set(asm_only)
    changed = True
    while changed:
        changed = False
        for fn, targets in refs.items():
            if fn not in tainted and targets & tainted:
                tainted.add(fn)
                changed = True
    return asm_only, tainted - asm_only


def rewrite_rtcd(cfgdir, tainted):
    """Take the assembler-dependent flavours out of the generated dispatch: a static binding
    `#define f f_sse2` becomes `#define f f_c`, a run-time assignment of a tainted flavour goes
    (the default assignment falls back to `_c`)."""
    n = 0
    for name in ("av1_rtcd.h", "aom_dsp_rtcd.h", "aom_scale_rtcd.h"):
        path = os.path.join(cfgdir, name)
        out = []
        for line in open(path + ".gen").read().splitlines():
            m = re.match(r"#define (\w+) (\w+)$", line)
            if m and m.group(2) in tainted:
                line = "#define %s %s_c" % (m.group(1), m.group(1))
                n += 1
            m = re.match(r"\s+(if \(flags & \w+\) )?(\w+) = (\w+);$", line)
            if m and m.group(3) in tainted:
                n += 1
                if m.group(1):
                    continue
                line = "    %s = %s_c;" % (m.group(2), m.group(2))
            out.append(line)
        write_if_changed(path, "\n".join(out) + "\n")
    return n


def build_flavour(flavour):
    fl = FLAVOURS[flavour]
    root = os.path.join(OUT, flavour)
    cfgdir = os.path.join(root, "config")
    os.makedirs(cfgdir, exist_ok=True)
    write_config(cfgdir, flavour)
    write_rtcd(cfgdir, fl["arch"])
    incs = ["-I" + REF, "-I" + root, "-I" + HERE, "-I" + os.path.join(REF, "apps"), "-I" + os.path.join(REF, "common")]
    x86 = []
    if flavour == "avx2":
        for d in X86_DIRS:
            for s in sorted(glob.glob(os.path.join(REF, d, "*.c"))):
                if os.path.relpath(s, REF) not in X86_EXCLUDE:
                    x86.append((s, isa_flag(s)))
        # plain-C statements of the assembler-only 1-D filter helpers (see the file's header)
        x86.append((os.path.join(HERE, "ref_asm_fallback.c"), []))
        for name in ("av1_rtcd.h", "aom_dsp_rtcd.h", "aom_scale_rtcd.h"):
            if not os.path.exists(os.path.join(cfgdir, name)):
                rewrite_rtcd(cfgdir, set())
        pass1, _failed = compile_all(x86, os.path.join(root, "obj_pass1"), incs, [], cfgdir)
        asm_only, tainted = asm_dependent(pass1)
        n = rewrite_rtcd(cfgdir, tainted)
        print("mk_ref[avx2]: %d assembler-only helpers; %d dependent wrappers, %d dispatch entries routed to _c"
              % (len(asm_only), len(tainted), n))
    else:
        rewrite_rtcd(cfgdir, set())
    srcs = []
    for d in LIB_DIRS:
        for s in sorted(glob.glob(os.path.join(REF, d, "*.c"))):
            if os.path.relpath(s, REF) not in EXCLUDE:
                srcs.append((s, []))
    srcs += x86
    srcs += [(os.path.join(REF, s), []) for s in LIB_EXTRA]
    srcs.append((os.path.join(cfgdir, "aom_config.c"), []))
    srcs.append((os.path.join(HERE, "ref_shim.c"), ["-DREFSHIM_DISPATCH"] if flavour == "avx2" else []))
    objs, failed = compile_all(srcs, os.path.join(root, "obj"), incs, [], cfgdir)
    # intrinsics files that need a configuration this build leaves off simply stay out
    hard = [f for f in failed if "/x86/" not in f[0]]
    if hard:
        for s, err in hard[:5]:
            sys.stderr.write("mk_ref: %s\n%s\n" % (s, err[-1500:]))
        raise SystemExit("mk_ref: %d source(s) failed to compile" % len(hard))
    for s, _e in failed:
        print("mk_ref[%s]: left out (does not compile in this configuration): %s" % (flavour, os.path.relpath(s, REF)))
    lib = os.path.join(OUT, "libaomref%s.so" % fl["suffix"])
    link = ["gcc", "-shared", "-o", lib] + objs + ["-fPIC", "-lm", "-lpthread"]
    subprocess.run(link, check=True)
    # undefined symbols = flavours that exist only as assembler (or in a left-out file): alias to _c
    undef = []
    for line in subprocess.run(["nm", "-D", lib], check=True, capture_output=True, text=True).stdout.splitlines():
        parts = line.split()
        if len(parts) == 2 and parts[0] == "U" and "@" not in parts[1] and not parts[1].startswith("_"):
            undef.append(parts[1])
    defined = set()
    for line in subprocess.run(["nm", "-D", "--defined-only", lib], check=True, capture_output=True,
                               text=True).stdout.splitlines():
        parts = line.split()
        if len(parts) == 3:
            defined.add(parts[2])
    syms, stubs = [], []
    for u in undef:
        c = re.sub(r"_(mmx|sse|sse2|sse3|ssse3|sse4_1|sse4_2|avx|avx2)$", "_c", u)
        if c != u and c in defined:
            syms.append("-Wl,--defsym=%s=%s" % (u, c))
        else:
            stubs.append(u)
    stub_c = os.path.join(root, "auto_stubs.c")
    with open(stub_c, "w") as f:
        f.write("/* abort() stubs for symbols of configurations this build leaves off */\n#include <stdio.h>\n#include <stdlib.h>\n")
        for u in stubs:
            f.write('void %s(void) { fprintf(stderr, "aomref stub called: %s\\n"); abort(); }\n' % (u, u))
    if syms or stubs:
        subprocess.run(link + syms + [stub_c], check=True)
    with open(os.path.join(root, "aliased_to_c.txt"), "w") as f:
        f.write("\n".join(sorted(s.split("=")[1] for s in syms)) + "\n")
    print("mk_ref[%s]: %d objects, %d asm-only flavours aliased to _c, %d stubs -> %s"
          % (flavour, len(objs), len(syms), len(stubs), lib))
    # the encoder CLI on top of the library
    app_objs, app_failed = compile_all([(os.path.join(REF, s), []) for s in APP_SOURCES],
                                       os.path.join(root, "obj_app"), incs, [], cfgdir)
    if app_failed:
        for s, err in app_failed[:3]:
            sys.stderr.write("mk_ref: %s\n%s\n" % (s, err[-1500:]))
        raise SystemExit("mk_ref: aomenc sources failed to compile")
    exe = os.path.join(OUT, "aomenc%s" % fl["suffix"])
    subprocess.run(["gcc", "-o", exe] + app_objs + [lib, "-Wl,-rpath,$ORIGIN", "-lm", "-lpthread"], check=True)
    print("mk_ref[%s]: built %s" % (flavour, exe))


def build_b200_install():
    """The `b200` RTCD flavour made executable (tools/gen_rtcd_b200.py): regenerated rtcd headers
    (type-checked against include/b200av1_rtcd.h by compiling them) and libb200install.so, the run-time
    binding for the already built avx2 reference library + aomenc_avx2."""
    root = os.path.join(OUT, "rtcd_b200")
    repo = os.path.dirname(HERE)
    prod = os.path.join(repo, "aom-av1-lavish_b200", "libb200av1.so")
    cfg = os.path.join(OUT, "avx2")
    subprocess.run([sys.executable, os.path.join(repo, "tools", "gen_rtcd_b200.py"), "--ref", REF, "--config",
                    os.path.join(cfg, "config"), "--out", root], check=True)
    incs = ["-I" + root, "-I" + REF, "-I" + cfg, "-I" + os.path.join(repo, "include")]
    for h in ("aom_dsp_rtcd", "av1_rtcd"):  # every _b200 prototype against the add_proto it replaces
        chk = os.path.join(root, "chk_%s.c" % h)
        with open(chk, "w") as f:
            f.write('#define RTCD_C\n#include "config/aom_config.h"\n#include "config_b200/%s.h"\n' % h)
        subprocess.run(["gcc", "-fsyntax-only", "-Werror", "-Wall", "-Wno-array-parameter", "-Wno-unused-function"]
                       + incs + [chk], check=True)
    if not os.path.exists(prod):
        print("mk_ref: %s not built yet; libb200install.so skipped" % prod)
        return
    lib = os.path.join(OUT, "libb200install.so")
    subprocess.run(["gcc", "-O2", "-fPIC", "-shared", "-Werror=incompatible-pointer-types", "-o", lib,
                    os.path.join(root, "b200_rtcd_install.c"), os.path.join(HERE, "b200_install_shim.c")] + incs +
                   [os.path.join(OUT, "libaomref_avx2.so"), prod, "-Wl,-rpath,$ORIGIN",
                    "-Wl,-rpath,$ORIGIN/../../aom-av1-lavish_b200"], check=True)
    print("mk_ref: built %s" % lib)


def main():
    if not os.path.isdir(REF):
        print("mk_ref: %s absent; keeping any prebuilt oracle/_ref" % REF)
        return 0
    os.makedirs(OUT, exist_ok=True)
    targets = [os.path.join(OUT, n) for n in ("libaomref.so", "libaomref_avx2.so", "aomenc", "aomenc_avx2")]
    install = os.path.join(OUT, "libb200install.so")
    repo = os.path.dirname(HERE)
    install_deps = [os.path.join(repo, "tools", "gen_rtcd_b200.py"), os.path.join(repo, "include", "b200av1_rtcd.h"),
                    os.path.join(HERE, "b200_install_shim.c"), __file__]
    newest = max(os.path.getmtime(__file__), os.path.getmtime(os.path.join(HERE, "ref_shim.c")),
                 os.path.getmtime(os.path.join(HERE, "av1o_me.h")),
                 os.path.getmtime(os.path.join(HERE, "ref_asm_fallback.c")))
    if all(os.path.exists(t) and os.path.getmtime(t) > newest for t in targets) and "--force" not in sys.argv:
        print("mk_ref: up to date")
    else:
        for flavour in ("generic", "avx2"):
            build_flavour(flavour)
    if (not os.path.exists(install) or "--force" in sys.argv or
            os.path.getmtime(install) < max([os.path.getmtime(p) for p in install_deps + targets[:2]])):
        build_b200_install()
    return 0


if __name__ == "__main__":
    sys.exit(main())
