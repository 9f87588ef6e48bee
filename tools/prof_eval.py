#!/usr/bin/env python3
"""Small driver for ncu: a few eval-mode fused forwards (grid-evaluation style: one scene, many points)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from deepsdf_b200.trainer import SceneTrainer  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 19
dev = torch.device("cuda:0")
tr = SceneTrainer(bench.SPECS, 8, dev, seed=0)
eng = tr.engine
ids = torch.tensor([3], device=dev)
cb = eng.scene_bias(tr.lat_vecs, ids)
ts = torch.zeros(N // 128, dtype=torch.int32, device=dev)
xyz = torch.rand(N, 3, device=dev) * 2 - 1
out = torch.empty(N, device=dev)
for _ in range(3):
    eng.decode_scenes(cb, ts, xyz, out=out)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5):
    eng.decode_scenes(cb, ts, xyz, out=out)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 5
print(f"eval forward N={N}: {ms:.3f} ms, {N / ms / 1e3:.1f} M rows/s, {N * 3.671e6 / ms / 1e9:.0f} algorithmic TFLOP/s, sdf mean {float(out.mean()):.4f}")
