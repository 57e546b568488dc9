"""The in-kernel guard (|residual| <= 4094 -> 3-instruction 32-bit butterflies) must be
backed by tools/range_proof.py for every transform size / 1-D type combination."""
import os
import re
import sys

import harness as H

sys.path.insert(0, os.path.join(H.ROOT, "tools"))


def test_guard_is_proved():
    import range_proof as R
    worst = 1e18
    for ts in range(19):
        W, Hh = R.TX_W[ts], R.TX_H[ts]
        kv = [0] if Hh == 64 else ([0, 2] if Hh == 32 else [0, 1, 2])
        kh = [0] if W == 64 else ([0, 2] if W == 32 else [0, 1, 2])
        for vk in kv:
            for hk in kh:
                worst = min(worst, R.analyse(ts, vk, hk))
    src = open(os.path.join(H.ROOT, "aom-av1-lavish_b200", "csrc", "txq_kernel.cuh")).read()
    guard = int(re.search(r"constexpr int kGuard = (\d+);", src).group(1))
    assert guard <= int(worst), (guard, worst)
    assert guard >= 1023  # any 10-bit residual always takes the fast path
