#!/usr/bin/env python3
"""dgrad chain time (1 M rows, L = 256) as a function of what it stashes (store_mask) and sums (sum_mask)."""
import ctypes as C
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from deepsdf_b200 import _lib as L  # noqa: E402
from deepsdf_b200.engine import FusedEngine  # noqa: E402
from deepsdf_b200.networks.deep_sdf_decoder import Decoder  # noqa: E402

dev = torch.device("cuda:0")
N = 1 << 20
torch.manual_seed(0)
dec = Decoder(256, dims=[512] * 8, geom_dimension=3, dropout=list(range(8)), dropout_prob=0.2, norm_layers=list(range(8)),
              latent_in=[4], weight_norm=True).to(dev).eval()
eng = FusedEngine(dec)
xyz = torch.rand(N, 3, device=dev) * 2 - 1
ts = torch.zeros(N // 128, dtype=torch.int32, device=dev)
cb = eng.scene_bias(torch.randn(1, 256, device=dev) * 0.1, torch.zeros(1, dtype=torch.int64, device=dev))
sdf = torch.empty(N, device=dev)
b = eng.grad_buffers(N)
io = L.TrainIO()
io.b_last = eng._b_last().data_ptr()
io.training, io.dropout_off, io.dy_unit = 1, 1, 1
io.dy_out, io.bitmask, io.n_rows_pad = b["dy"].data_ptr(), b["bitmask"].data_ptr(), N
st = eng._st()
L.check(eng.lib.dsdf_fused_forward(C.byref(eng.desc), eng.pack.data_ptr(), cb.data_ptr(), ts.data_ptr(), xyz.data_ptr(), 3, N, sdf.data_ptr(), C.byref(io), st))
ALL, TRAIN = 0xFFFFFFFF, 0xFFFFFFFE
cfgs = (("store all, sums all", TRAIN, ALL), ("store all, no sums", TRAIN, 0), ("no store, sums all", 0, ALL), ("no store, no sums", 0, 0))
best = {}
for rnd in range(4):
    for name, store, sums in cfgs:
        f = lambda: eng.lib.dsdf_fused_dgrad(C.byref(eng.desc), eng.pack.data_ptr(), C.byref(eng._tb), xyz.data_ptr(), 3, N, 0, store, sums, st)
        L.check(f())
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            L.check(f())
        e1.record()
        torch.cuda.synchronize()
        best.setdefault(name, []).append(round(e0.elapsed_time(e1) / 10, 3))
for k, v in best.items():
    print(f"{k:32s} {v}")
