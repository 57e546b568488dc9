#!/usr/bin/env python3
"""Dev experiment (GPU box): time txq_kernel variants per TX size to see what the config[3]
path (pixel input, highbd quantiser, block error) costs over the config[1] path."""
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench as B  # noqa: E402
import b200av1  # noqa: E402

dev = torch.device("cuda", 0)
ctx = b200av1.Context(0)
stream = torch.cuda.Stream(device=dev)
planes = [(3840, 2160), (1920, 1080), (1920, 1080)]
fe = sum(w * h for w, h in planes)
nf = 3
rng = np.random.default_rng(1)
src = rng.integers(300, 700, nf * fe).astype(np.uint16)
pred = (src.astype(np.int32) + rng.integers(-40, 41, nf * fe)).astype(np.uint16)
resid = (src.astype(np.int32) - pred.astype(np.int32)).astype(np.int16)
d_src = torch.from_numpy(src.view(np.int16)).to(dev)
d_pred = torch.from_numpy(pred.view(np.int16)).to(dev)
d_res = torch.from_numpy(resid).to(dev)
qrows = B.build_qrows(10)
d_q = torch.from_numpy(qrows.view(np.uint8)).to(dev)
n_out = nf * fe
d_c, d_qc, d_dq = (torch.empty(n_out, dtype=torch.int32, device=dev) for _ in range(3))
variants = [("resid fp", 0, b200av1.Q_FP, 0), ("resid fp +err", 0, b200av1.Q_FP, 1),
            ("resid fp_hbd +err", 0, b200av1.Q_FP_HBD, 1), ("pix16 fp_hbd", 2, b200av1.Q_FP_HBD, 0),
            ("pix16 fp_hbd +err", 2, b200av1.Q_FP_HBD, 1)]
sizes = [int(s) for s in sys.argv[1].split(",")] if len(sys.argv) > 1 else list(range(19))
print("%-8s" % "tx" + "".join("%20s" % v[0] for v in variants))
tot = np.zeros(len(variants))
for ts in sizes:
    blocks, samples = B.make_descriptors(ts, nf, planes, fe, 5)
    preds = np.zeros(blocks.size, b200av1.TXPRED_DTYPE)
    preds["pred_offset"], preds["pred_stride"] = blocks["src_offset"], blocks["src_stride"]
    d_b = torch.from_numpy(blocks.view(np.uint8)).to(dev)
    d_p = torch.from_numpy(preds.view(np.uint8)).to(dev)
    d_eob = torch.zeros(blocks.size, dtype=torch.int16, device=dev)
    d_dist = torch.zeros(blocks.size * 2, dtype=torch.int64, device=dev)
    line = "%-8s" % ("%dx%d" % (B.TX_W[ts], B.TX_H[ts]))
    for vi, (name, in_kind, qk, err) in enumerate(variants):
        l = b200av1.TxqLaunch()
        l.tx_size, l.in_kind, l.q_kind, l.bd = ts, in_kind, qk, 10
        l.flags = b200av1.TXQ_STORE_COEFF | (b200av1.TXQ_BLOCK_ERROR if err else 0)
        l.n_blocks = int(blocks.size)
        l.blocks, l.preds = d_b.data_ptr(), d_p.data_ptr()
        l.src = d_res.data_ptr() if in_kind == 0 else d_src.data_ptr()
        l.pred, l.qrows = d_pred.data_ptr(), d_q.data_ptr()
        l.coeff, l.qcoeff, l.dqcoeff = d_c.data_ptr(), d_qc.data_ptr(), d_dq.data_ptr()
        l.eob, l.dist = d_eob.data_ptr(), d_dist.data_ptr()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        ms = 0.0
        for it in range(6):
            ev[0].record(stream)
            ctx.txq_launch(l, stream.cuda_stream)
            ev[1].record(stream)
            torch.cuda.synchronize(dev)
            if it >= 2:
                ms += ev[0].elapsed_time(ev[1]) / 4
        nbytes = blocks.size * (B.TX_W[ts] * B.TX_H[ts] * (2 if in_kind == 0 else 4) + B.max_eob(ts) * 12 + 2 + 16 * err)
        line += "%12.3f ms %5.0f" % (ms, nbytes / ms / 1e6)
        tot[vi] += ms
    print(line)
print("%-8s" % "total" + "".join("%17.3f ms" % t for t in tot))
