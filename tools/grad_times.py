#!/usr/bin/env python3
"""Per-kernel CUDA-event times of the input-gradient path (forward in gates mode, dgrad chain, input_grad GEMM) on a
2 M-row grid chunk, L = 256 and L = 2."""
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
N = 1 << 21
for Lz in (256, 2):
    torch.manual_seed(0)
    dec = Decoder(Lz, dims=[512] * 8, geom_dimension=3, dropout=list(range(8)), dropout_prob=0.2, norm_layers=list(range(8)),
                  latent_in=[4], weight_norm=True).to(dev).eval()
    eng = FusedEngine(dec)
    xyz = torch.rand(N, 3, device=dev) * 2 - 1
    ts = torch.zeros(N // 128, dtype=torch.int32, device=dev)
    cb = eng.scene_bias(torch.randn(1, Lz, device=dev) * 0.1, torch.zeros(1, dtype=torch.int64, device=dev))
    G = torch.empty(N, Lz + 3, device=dev)
    sdf = torch.empty(N, device=dev)
    b = eng.grad_buffers(N)
    io = L.TrainIO()
    io.b_last = eng._b_last().data_ptr()
    io.training, io.dropout_off, io.dy_unit = 1, 1, 1
    io.dy_out, io.bitmask, io.n_rows_pad = b["dy"].data_ptr(), b["bitmask"].data_ptr(), N
    lat = eng.lat_layer
    store = 1 | (1 << lat)
    st = eng._st()
    calls = [
        ("forward(gates)", lambda: eng.lib.dsdf_fused_forward(C.byref(eng.desc), eng.pack.data_ptr(), cb.data_ptr(), ts.data_ptr(), xyz.data_ptr(), 3, N, sdf.data_ptr(), C.byref(io), st)),
        ("dgrad(store 2)", lambda: eng.lib.dsdf_fused_dgrad(C.byref(eng.desc), eng.pack.data_ptr(), C.byref(eng._tb), xyz.data_ptr(), 3, N, 0, store, 0, st)),
        ("input_grad", lambda: eng.lib.dsdf_fused_input_grad(C.byref(eng.desc), eng.pack.data_ptr(), C.byref(eng._tb), N, G.data_ptr(), G.stride(0), st)),
    ]
    for _ in range(2):
        for _, f in calls:
            L.check(f())
    torch.cuda.synchronize()
    out = {}
    for name, f in calls:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            L.check(f())
        e1.record()
        torch.cuda.synchronize()
        out[name] = round(e0.elapsed_time(e1) / 5, 3)
    print(f"L={Lz}: ms per 2 M rows", out)
    del eng, G, b
    torch.cuda.empty_cache()
