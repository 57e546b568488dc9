"""The `b200` RTCD flavour, executable:
* CPU: tools/gen_rtcd_b200.py runs the reference's own rtcd.pl (patched copy, never the reference
  tree) on its *_rtcd_defs.pl with `b200` specialisations; every drop-in becomes a pointer bound by
  `if (b200_device_count() > 0) fn = fn_b200;`, and the generated headers compile together with
  include/b200av1_rtcd.h, i.e. all `_b200` prototypes are the add_proto types they replace.
* GPU: the UNMODIFIED aomenc_avx2 (reference build, run-time dispatch) with the transform / quantiser /
  block-error / inverse-transform groups bound to `_b200` by oracle/_ref/libb200install.so produces a
  byte-identical .ivf."""
import os
import re
import subprocess
import sys

import pytest

import harness as H

REF = "/root/reference"
OUT = os.path.join(H.ORACLE_DIR, "_ref")


@pytest.mark.skipif(not os.path.isdir(REF) or not os.path.exists(os.path.join(OUT, "avx2", "config", "aom_config.h")),
                    reason="reference tree / avx2 reference build absent")
def test_generated_rtcd_binds_every_dropin(tmp_path):
    r = subprocess.run([sys.executable, os.path.join(H.ROOT, "tools", "gen_rtcd_b200.py"), "--ref", REF, "--config",
                        os.path.join(OUT, "avx2", "config"), "--out", str(tmp_path)], check=True,
                       capture_output=True, text=True)
    m = re.search(r"(\d+) drop-ins; stock headers: (\d+) run-time pointers, (\d+) static, (\d+) not RTCD .*bind (\d+)",
                  r.stdout)
    n, ptr, static, other, bound = map(int, m.groups())
    assert n > 1000 and bound == n - other and other <= 1 and ptr + static + other == n
    gen = open(tmp_path / "config_b200" / "aom_dsp_rtcd.h").read() + open(tmp_path / "config_b200" / "av1_rtcd.h").read()
    for fn in ("av1_fwd_txfm2d_16x16", "av1_quantize_fp", "av1_highbd_quantize_fp", "av1_block_error",
               "aom_sad64x64x4d", "aom_sub_pixel_variance64x64", "aom_masked_sad16x16", "aom_obmc_variance8x8",
               "av1_inv_txfm_add", "aom_hadamard_16x16"):
        assert "RTCD_EXTERN" in gen and re.search(r"\(\*%s\)\(" % fn, gen), fn
        assert "if (b200_device_count() > 0) %s = %s_b200;" % (fn, fn) in gen, fn
    incs = ["-I" + str(tmp_path), "-I" + REF, "-I" + os.path.join(OUT, "avx2"), "-I" + os.path.join(H.ROOT, "include")]
    for h in ("aom_dsp_rtcd", "av1_rtcd"):
        chk = tmp_path / ("chk_%s.c" % h)
        chk.write_text('#define RTCD_C\n#include "config/aom_config.h"\n#include "config_b200/%s.h"\n' % h)
        subprocess.run(["gcc", "-fsyntax-only", "-Werror", "-Wall", "-Wno-array-parameter", "-Wno-unused-function"] +
                       incs + [str(chk)], check=True)
    inst = open(tmp_path / "b200_rtcd_install.c").read()
    assert inst.count("_b200;") == ptr


@pytest.mark.gpu
@pytest.mark.skipif(not (os.path.exists(os.path.join(OUT, "libb200install.so")) and
                         os.path.exists(os.path.join(OUT, "aomenc_avx2")) and H.host_has_avx2()),
                    reason="libb200install.so / aomenc_avx2 not built or no AVX2 host")
def test_aomenc_with_b200_flavour_is_byte_identical(tmp_path):
    sys.path.insert(0, os.path.join(H.ROOT, "baseline"))
    import gen_y4m
    clip = str(tmp_path / "qcif.y4m")
    gen_y4m.write_y4m(clip, 176, 144, 3)
    exe = os.path.join(OUT, "aomenc_avx2")
    args = [exe, "--cpu-used=6", "--good", "--threads=1", "--limit=3", "-q"]
    subprocess.run(args + ["-o", str(tmp_path / "stock.ivf"), clip], check=True, capture_output=True, timeout=300)
    env = dict(os.environ, LD_PRELOAD=os.path.join(OUT, "libb200install.so"), B200_RTCD_GROUPS="15")
    r = subprocess.run(args + ["-o", str(tmp_path / "b200.ivf"), clip], env=env, capture_output=True, timeout=1500)
    assert r.returncode == 0, r.stderr[-800:]
    m = re.search(rb"(\d+) RTCD symbols bound", r.stderr)
    assert m and int(m.group(1)) >= 40, r.stderr[-400:]
    a, b = open(tmp_path / "stock.ivf", "rb").read(), open(tmp_path / "b200.ivf", "rb").read()
    assert len(a) > 500 and a == b
