#!/usr/bin/env python3
"""range_proof.py -- worst-case magnitude analysis of the AV1 forward transform networks.

Purpose: the CUDA kernels evaluate half_btf (av1/common/av1_txfm.h:80) as
(w0*x0 + w1*x1 + round) >> bit in wrapping 32-bit arithmetic (3 instructions) whenever
the block passes a conformance guard |residual| <= GUARD, and fall back to the literal
64-bit form otherwise.  The two forms agree iff the true value of w0*x0 + w1*x1 + round
fits in int32.  This script PROVES a safe GUARD: every node of every 1-D network is
carried as a linear form over the N inputs (exact float coefficients) plus an
accumulated rounding-error bound; for inputs in [-M, M]^N the largest possible
|w0*x0 + w1*x1| at a rotation node is M * L1norm(w0*g0 + w1*g1) + error terms, attained
by the sign pattern of the coefficients.  The column pass feeds the row pass with
values bounded by its own worst-case output.  Output: per (tx_size, column kind, row
kind) the largest M for which no rotation sum, no rounding add and no sqrt(2) rescale
can leave int32.  tests/test_range_proof.py asserts min M >= 4095 (= any 12-bit residual).
"""
import math
import sys

import numpy as np

TX_W = [4, 8, 16, 32, 64, 4, 8, 8, 16, 16, 32, 32, 64, 4, 16, 8, 32, 16, 64]
TX_H = [4, 8, 16, 32, 64, 8, 4, 16, 8, 32, 16, 64, 32, 16, 4, 32, 8, 64, 16]
SHIFT = [(2, 0, 0), (2, -1, 0), (2, -2, 0), (2, -4, 0), (0, -2, -2), (2, -1, 0), (2, -1, 0),
         (2, -2, 0), (2, -2, 0), (2, -4, 0), (2, -4, 0), (0, -2, -2), (2, -4, -2), (2, -1, 0),
         (2, -1, 0), (2, -2, 0), (2, -2, 0), (0, -2, 0), (2, -4, 0)]
COS_COL = [[13, 13, 13, 0, 0], [13, 13, 13, 12, 0], [13, 13, 13, 12, 13], [0, 13, 13, 12, 13],
           [0, 0, 13, 12, 13]]
COS_ROW = [[13, 13, 12, 0, 0], [13, 13, 13, 12, 0], [13, 13, 12, 13, 12], [0, 12, 13, 12, 11],
           [0, 0, 12, 11, 10]]
SINPI13 = [0, 2642, 4964, 6689, 7606]


def cospi(bit):
    return [int(math.floor(math.cos(math.pi * j / 128) * (1 << bit) + 0.5)) for j in range(64)]


def bitrev(v, bits):
    r = 0
    for i in range(bits):
        r |= ((v >> i) & 1) << (bits - 1 - i)
    return r


class Node:
    """linear form g over the inputs + rounding error bound e"""
    __slots__ = ("g", "e")

    def __init__(self, g, e=0.0):
        self.g, self.e = g, e

    def __add__(self, o):
        return Node(self.g + o.g, self.e + o.e)

    def __sub__(self, o):
        return Node(self.g - o.g, self.e + o.e)

    def neg(self):
        return Node(-self.g, self.e)


class Tracker:
    def __init__(self):
        self.worst_l1 = 0.0   # max over nodes of L1 of the pre-shift sum, per unit M
        self.worst_add = 0.0  # additive part (errors + rounding constant)

    def note(self, l1, add):
        # keep the pair that is worst for large M (dominant term) but remember max add
        if l1 > self.worst_l1:
            self.worst_l1 = l1
        if add > self.worst_add:
            self.worst_add = add


def rot(tr, c, bit, s0, a0, x0, s1, a1, x1):
    w0, w1 = s0 * c[a0], s1 * c[a1]
    g = w0 * x0.g + w1 * x1.g
    tr.note(np.abs(g).sum(), abs(w0) * x0.e + abs(w1) * x1.e + (1 << (bit - 1)))
    return Node(g / (1 << bit), (abs(w0) * x0.e + abs(w1) * x1.e) / (1 << bit) + 0.5)


def dct_odd(tr, u, M, c, bit):
    p = M
    while p >= 4:
        for j in range(M // 2):
            t, k = j % p, j // p
            if p == M:
                A = 32
            else:
                n = M // (2 * p)
                A = 16 // n + (64 // n) * bitrev(k, n.bit_length() - 1)
            lo, hi = u[j], u[M - 1 - j]
            if p // 4 <= t < p // 2:
                u[j] = rot(tr, c, bit, -1, A, lo, +1, 64 - A, hi)
                u[M - 1 - j] = rot(tr, c, bit, +1, A, hi, +1, 64 - A, lo)
            elif p // 2 <= t < 3 * p // 4:
                u[j] = rot(tr, c, bit, -1, 64 - A, lo, -1, A, hi)
                u[M - 1 - j] = rot(tr, c, bit, +1, 64 - A, hi, -1, A, lo)
        G = p // 2
        for g in range(M // G):
            for i in range(G // 2):
                ia, ib = g * G + i, g * G + G - 1 - i
                a, b = u[ia], u[ib]
                if g % 2 == 0:
                    u[ia], u[ib] = a + b, a - b
                else:
                    u[ia], u[ib] = b - a, b + a
        p //= 2
    lb = M.bit_length() - 2
    for j in range(M // 2):
        Y = 32 // M + (128 // M) * bitrev(j, lb)
        lo, hi = u[j], u[M - 1 - j]
        u[j] = rot(tr, c, bit, +1, 64 - Y, lo, +1, Y, hi)
        u[M - 1 - j] = rot(tr, c, bit, +1, 64 - Y, hi, -1, Y, lo)


def dct_core(tr, v, c, bit):
    n = len(v)
    if n == 2:
        a, b = v
        return [rot(tr, c, bit, +1, 32, a, +1, 32, b), rot(tr, c, bit, -1, 32, b, +1, 32, a)]
    m = n // 2
    s = [v[i] + v[n - 1 - i] for i in range(m)]
    d = [v[m - 1 - j] - v[m + j] for j in range(m)]
    ev = dct_core(tr, s, c, bit)
    dct_odd(tr, d, m, c, bit)
    return ev + d


def fdct(tr, x, bit):
    n = len(x)
    v = dct_core(tr, list(x), cospi(bit), bit)
    out = [None] * n
    for i in range(n):
        out[bitrev(i, n.bit_length() - 1)] = v[i]
    return out


def fadst(tr, x, bit):
    n = len(x)
    c = cospi(bit)
    if n == 4:
        s = [v * (1 << (bit - 13)) if bit >= 13 else v / (1 << (13 - bit)) for v in SINPI13]
        x0, x1, x2, x3 = x
        a7 = x0 + x1 - x3
        forms = [x0.g * s[1] + x1.g * s[2] + x3.g * s[4] + x2.g * s[3],
                 a7.g * s[3],
                 x0.g * s[4] - x1.g * s[1] + x3.g * s[2] - x2.g * s[3],
                 (x0.g * s[4] - x1.g * s[1] + x3.g * s[2]) - (x0.g * s[1] + x1.g * s[2] + x3.g * s[4]) + x2.g * s[3]]
        out = []
        for g in forms:
            tr.note(np.abs(g).sum(), 8 * 8192 * max(v.e for v in x) + (1 << (bit - 1)))
            out.append(Node(g / (1 << bit), 8 * max(v.e for v in x) + 0.5))
        return out
    perm = {8: [0, -7, -3, 4, -1, 6, 2, -5],
            16: [0, -15, -7, 8, -3, 12, 4, -11, -1, 14, 6, -9, 2, -13, -5, 10]}[n]
    v = [x[-p].neg() if p < 0 else x[p] for p in perm]
    g = 4
    while g <= n:
        pairs = g // 4
        for base in range(0, n, g):
            for q in range(pairs):
                i0 = base + g // 2 + 2 * q
                a, b = v[i0], v[i0 + 1]
                if pairs == 1:
                    v[i0] = rot(tr, c, bit, +1, 32, a, +1, 32, b)
                    v[i0 + 1] = rot(tr, c, bit, +1, 32, a, -1, 32, b)
                else:
                    half = pairs // 2
                    A = 16 // half + (64 // half) * (q % half)
                    if q < half:
                        v[i0] = rot(tr, c, bit, +1, A, a, +1, 64 - A, b)
                        v[i0 + 1] = rot(tr, c, bit, +1, 64 - A, a, -1, A, b)
                    else:
                        v[i0] = rot(tr, c, bit, -1, 64 - A, a, +1, A, b)
                        v[i0 + 1] = rot(tr, c, bit, +1, A, a, +1, 64 - A, b)
            for i in range(g // 2):
                a, b = v[base + i], v[base + g // 2 + i]
                v[base + i], v[base + g // 2 + i] = a + b, a - b
        g *= 2
    for i in range(n // 2):
        A = 32 // n + (128 // n) * i
        a, b = v[2 * i], v[2 * i + 1]
        v[2 * i] = rot(tr, c, bit, +1, A, a, +1, 64 - A, b)
        v[2 * i + 1] = rot(tr, c, bit, +1, 64 - A, a, -1, A, b)
    out = [None] * n
    for i in range(n // 2):
        out[2 * i] = v[2 * i + 1]
        out[2 * i + 1] = v[n - 2 - 2 * i]
    return out


def fidentity(tr, x):
    n = len(x)
    mul = {4: 5793 / 4096, 8: 2.0, 16: 2 * 5793 / 4096, 32: 4.0}[n]
    out = []
    for v in x:
        # the fast form computes x*1697 (or x*3394) + 2048 in 32 bits
        if n in (4, 16):
            k = 1697 if n == 4 else 3394
            tr.note(np.abs(v.g).sum() * k, k * v.e + 2048)
        out.append(Node(v.g * mul, v.e * mul + 0.5))
    return out


def txfm1d(tr, kind, x, bit):
    if kind == 0:
        return fdct(tr, x, bit)
    if kind == 1:
        return fadst(tr, x, bit)
    return fidentity(tr, x)


def out_bound(nodes):
    """(L1 per unit M, additive) of the largest-magnitude output"""
    return max(np.abs(v.g).sum() for v in nodes), max(v.e for v in nodes)


def analyse(ts, vkind, hkind):
    W, H = TX_W[ts], TX_H[ts]
    wi, hi = W.bit_length() - 3, H.bit_length() - 3
    up, d1, d2 = SHIFT[ts][0], -SHIFT[ts][1], -SHIFT[ts][2]
    # column pass: inputs = residual << up, |residual| <= M
    trc = Tracker()
    x = [Node(np.eye(H)[i] * (1 << up)) for i in range(H)]
    y = txfm1d(trc, vkind, x, COS_COL[wi][hi])
    l1c, ec = out_bound(y)
    # rounding add (x + 2^(d1-1)) must fit as well: negligible next to rotation sums
    # row pass: every input is an independent column-pass output, |.| <= l1c*M/2^d1 + ec + 1
    trr = Tracker()
    x2 = [Node(np.eye(W)[i]) for i in range(W)]
    y2 = txfm1d(trr, hkind, x2, COS_ROW[wi][hi])
    l1r, er = out_bound(y2)
    lim = float((1 << 31) - 1)
    # solve for M: column nodes: trc.worst_l1 * M + trc.worst_add <= lim
    m_col = (lim - trc.worst_add) / trc.worst_l1 if trc.worst_l1 else float("inf")
    # row nodes with input bound B = l1c*M/2^d1 + ec/2^d1 + 1: trr.worst_l1 * B + add <= lim
    if trr.worst_l1:
        b_max = (lim - trr.worst_add) / trr.worst_l1
        m_row = (b_max - 1 - ec / (1 << d1)) * (1 << d1) / l1c
    else:
        m_row = float("inf")
    # final rescale of 2:1 rectangles: v*1697 + 2048 in 32 bits, v <= l1r*B/2^d2 + ..
    m_rect = float("inf")
    if W == 2 * H or H == 2 * W:
        vmax = (lim - 2048) / 1697
        b_ok = (vmax - 1 - er) * (1 << d2) / l1r
        m_rect = (b_ok - 1 - ec / (1 << d1)) * (1 << d1) / l1c
    return min(m_col, m_row, m_rect)


def main():
    worst = float("inf")
    for ts in range(19):
        W, H = TX_W[ts], TX_H[ts]
        kinds_v = [0] if H == 64 else ([0, 2] if H == 32 else [0, 1, 2])
        kinds_h = [0] if W == 64 else ([0, 2] if W == 32 else [0, 1, 2])
        row = []
        for vk in kinds_v:
            for hk in kinds_h:
                m = analyse(ts, vk, hk)
                row.append(m)
                worst = min(worst, m)
        print("%-6s min safe |residual| = %10.0f" % ("%dx%d" % (W, H), min(row)))
    print("overall: fast 32-bit butterflies are exact for |residual| <= %d" % int(worst))
    return int(worst)


if __name__ == "__main__":
    main()
