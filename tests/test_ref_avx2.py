"""The SIMD flavour of the reference (oracle/_ref/libaomref_avx2.so, built by oracle/mk_ref.py with the
x86 intrinsics and run-time dispatch) against its generic C flavour on the bench workload: the timed
`reference-avx2` CPU baseline must compute exactly what the `_c` path -- the parity oracle -- computes."""
import ctypes as C
import os
import sys

import numpy as np
import pytest

import harness as H

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

pytestmark = pytest.mark.skipif(H.ref() is None or H.ref_avx2() is None,
                                reason="oracle/_ref flavours not built or host without AVX2")


def test_flavours_identify_themselves():
    assert H.ref().refshim_simd() == 0
    assert H.ref_avx2().refshim_simd() == 1


@pytest.mark.parametrize("sizes", [[0, 1, 5, 13], [2, 7, 9, 16], [3, 4, 11, 18]])
def test_txq_sweep_avx2_equals_c(sizes):
    import bench
    frames, planes, _luma = bench.make_residual_frames(1, 3)
    # a band of the frame is enough: cut the descriptor list
    outs = []
    for R in (H.ref(), H.ref_avx2()):
        _t, _s, o = bench.cpu_run_sweep(H, H.oracle(), R, frames, planes, frames.shape[1], 3, 4, sizes,
                                        keep_outputs=True)
        outs.append(o)
    for ts in sizes:
        for a, b, name in zip(outs[0][ts][1:], outs[1][ts][1:], ("coeff", "qcoeff", "dqcoeff", "eob")):
            assert np.array_equal(a, b), "tx_size %d: %s differs between the _c and the SIMD flavour" % (ts, name)


def test_motion_search_avx2_equals_c():
    import bench
    _frames, _planes, luma = bench.make_residual_frames(1, 5)
    work = bench.MeWork(luma, [0], 5)
    res = [bench.cpu_run_me(H, H.oracle(), R, work, 0, 4) for R in (H.ref(), H.ref_avx2())]
    for i in range(work.per_frame):
        assert res[0][1][i].key() == res[1][1][i].key(), "full-pel search %d" % i
        assert res[0][2][i].key(True) == res[1][2][i].key(True), "sub-pel search %d" % i
