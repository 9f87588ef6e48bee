#!/usr/bin/env python3
"""Small driver for ncu: the split-fp16 high-precision eval forward (chain_exact_kernel) over 2^20 grid points of one code."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from deepsdf_b200.networks.deep_sdf_decoder import Decoder  # noqa: E402
from deepsdf_b200.utils import fused_engine_for  # noqa: E402

dev = torch.device("cuda:0")
torch.manual_seed(0)
dec = Decoder(bench.SPECS["CodeLength"], **bench.SPECS["NetworkSpecs"]).to(dev).eval()
eng = fused_engine_for(dec)
n = 1 << 20
lat = (torch.randn(1, bench.SPECS["CodeLength"], device=dev) / 16).contiguous()
zero = torch.zeros(1, dtype=torch.int64, device=dev)
cbx = eng.scene_bias_exact(lat, zero)
xyz = (torch.rand(n, 3, device=dev) * 2 - 1).contiguous()
ts = torch.zeros(n // 128, dtype=torch.int32, device=dev)
out = torch.empty(n, device=dev)
for _ in range(2):
    eng.decode_scenes(cbx, ts, xyz, out=out, exact=True)
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
eng.decode_scenes(cbx, ts, xyz, out=out, exact=True)
b.record()
torch.cuda.synchronize()
print(f"exact forward N={n}: {a.elapsed_time(b):.3f} ms, {n / a.elapsed_time(b) / 1e3:.1f} M rows/s, sdf mean {float(out.mean()):.4f}")
