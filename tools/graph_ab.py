#!/usr/bin/env python3
"""Eager launches vs CUDA-graph replay of the same training step, alternating in one process (power / thermal drift
cancels): ms per step at 64, 32, 16 and 8 scenes x 16384 rows."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from deepsdf_b200.trainer import SceneTrainer  # noqa: E402

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 20
dev = torch.device("cuda:0")
tr = SceneTrainer(bench.SPECS, bench.SCENES_TOTAL, dev, seed=0)
gen = torch.Generator(device=dev).manual_seed(2)
ids = torch.arange(64) * 3
batch = bench.make_batch(gen, dev, ids, bench.SCENES_TOTAL)
ids_d = ids.to(dev)
g_rows = 64 * bench.SAMPLES_PER_SCENE
for s in (64, 32, 16, 8):
    b, i = batch[:s * bench.SAMPLES_PER_SCENE], ids_d[:s].contiguous()
    res = {"eager": [], "graph": []}
    for rep in range(3):
        for mode in ("graph", "eager"):
            fn = tr.step if mode == "eager" else tr.step_graphed
            for _ in range(3):
                fn(b, i, epoch=1, global_rows=g_rows)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(steps):
                fn(b, i, epoch=1, global_rows=g_rows)
            e1.record()
            torch.cuda.synchronize()
            res[mode].append(round(e0.elapsed_time(e1) / steps, 4))
    print(s, "scenes:", res)
