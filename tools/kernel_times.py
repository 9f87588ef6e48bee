#!/usr/bin/env python3
"""Per-kernel CUDA-event times of the fused training step (config #2), for A/B experiments: prints the mean ms of
forward / dgrad / wgrad / finish over a few steps."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from deepsdf_b200.trainer import SceneTrainer  # noqa: E402

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 10
dev = torch.device("cuda:0")
tr = SceneTrainer(bench.SPECS, bench.SCENES_TOTAL, dev, seed=0)
gen = torch.Generator(device=dev).manual_seed(2)
ids = torch.arange(64) * 3
batch = bench.make_batch(gen, dev, ids, bench.SCENES_TOTAL)
ids_d = ids.to(dev)
for _ in range(3):
    tr.step(batch, ids_d, epoch=1)
torch.cuda.synchronize()
tr.engine.kernel_events = []
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(steps):
    loss = tr.step(batch, ids_d, epoch=1)
e1.record()
torch.cuda.synchronize()
per = {}
for name, a, b in tr.engine.kernel_events:
    per.setdefault(name, []).append(a.elapsed_time(b))
print({k: round(sum(v) / len(v), 3) for k, v in per.items()}, "step ms", round(e0.elapsed_time(e1) / steps, 3), "loss", float(loss[0]))
