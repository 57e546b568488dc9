#!/usr/bin/env python3
"""Pick shared-memory pitches for the txq v4 kernel: for every transform size and its
(CPL, RPL) lane mapping, search the transpose-tile row pitch LDX, slot pitch XS and the
staging slot pitch so that column loads, column stores and row loads are free of bank
conflicts.  Prints a table that txq4_kernel.cuh mirrors in constexpr form."""
import itertools, sys

def wavefronts(addrs_bytes, size):
    """addrs_bytes: per-lane byte address (32 lanes), size: access bytes. Returns worst
    number of wavefronts per phase relative to the ideal (1)."""
    lanes_per_phase = 32 if size <= 4 else (16 if size == 8 else 8)
    worst = 1
    for p in range(0, 32, lanes_per_phase):
        banks = {}
        for l in range(p, p + lanes_per_phase):
            a = addrs_bytes[l]
            for w in range(a // 4, (a + max(size, 1) + 3) // 4 if size >= 4 else a // 4 + 1):
                banks.setdefault(w % 32, set()).add(w)
        worst = max(worst, max(len(s) for s in banks.values()))
    return worst

def cfg(W, H):
    Hk, Wk = min(H, 32), min(W, 32)
    table = {(8, 8): (2, 2), (4, 16): (1, 4), (16, 4): (4, 1), (8, 16): (2, 4), (16, 8): (4, 2),
             (16, 16): (2, 2), (8, 32): (1, 4), (32, 8): (4, 1), (16, 32): (1, 2), (32, 16): (2, 1),
             (4, 4): (1, 1), (4, 8): (1, 2), (8, 4): (2, 1)}
    CPL, RPL = table.get((W, H), (1, 1))
    LC, LR = W // CPL, Hk // RPL
    TB = max(1, 32 // min(LC, LR))
    return CPL, RPL, LC, LR, TB, Hk

def check(W, H, esz, LDX, XS, SLOT):
    CPL, RPL, LC, LR, TB, Hk = cfg(W, H)
    res = []
    # column load (round 0 is representative when CI > 1: groups shift by a constant)
    for r in (0, 1, H - 1):
        a = []
        for lane in range(32):
            slot, g = lane % TB, (lane // TB) % LC
            a.append(slot * SLOT + r * W * esz + g * CPL * esz)
        res.append(wavefronts(a, CPL * esz))
    cl = max(res)
    res = []
    for r in (0, 1, Hk - 1):
        a = []
        for lane in range(32):
            slot, g = lane % TB, (lane // TB) % LC
            a.append(4 * (slot * XS + r * LDX + g * CPL))
        res.append(wavefronts(a, 4 * CPL))
    cs = max(res)
    res = []
    for j in range(RPL):
        for k in range(0, W, 4):
            a = []
            for lane in range(32):
                slot, g = lane % TB, (lane // TB) % LR
                a.append(4 * (slot * XS + (g * RPL + j) * LDX + k))
            res.append(wavefronts(a, 16 if W >= 4 else 4 * W))
    rl = max(res)
    return cl, cs, rl

sizes = [(8, 8), (4, 16), (16, 4), (8, 16), (16, 8), (16, 16), (8, 32), (32, 8), (16, 32), (32, 16),
         (32, 32), (4, 8), (8, 4), (4, 4)]
for (W, H) in sizes:
    CPL, RPL, LC, LR, TB, Hk = cfg(W, H)
    best = None
    for esz in (2,):
        tile = H * W * esz
        for LDX in range(W, W + 17, 4):
            if CPL > 1 and LDX % CPL: continue
            for XS in range(Hk * LDX, Hk * LDX + 68, 4):
                for SLOT in range((tile + 15) // 16 * 16, tile + 16 * 12, 16 if W * esz >= 16 else W * esz):
                    c = check(W, H, esz, LDX, XS, SLOT)
                    cost = (sum(c), LDX, XS, SLOT)
                    if best is None or cost < best[0]:
                        best = (cost, c)
    (s, LDX, XS, SLOT), c = best
    print(f"{W:2d}x{H:<2d} CPL {CPL} RPL {RPL} LC {LC:2d} LR {LR:2d} TB {TB:2d}  LDX {LDX:3d} (W+{LDX-W}) XS {XS:5d} (+{XS-Hk*LDX:2d}) SLOT {SLOT:5d} (+{SLOT-H*W*2:3d})  wavefronts col-ld {c[0]} col-st {c[1]} row-ld {c[2]}")
