#!/usr/bin/env python3
"""Step time of the per-rank shards of a strong-scaling run (64 / n scenes x 16384 rows), alone on one GPU:
   python tools/shard_times.py [steps]      prints ms/step for 64, 32, 16, 8 scenes (DSDF_NO_SIDE_STREAM=1 for the A/B)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from deepsdf_b200.trainer import SceneTrainer  # noqa: E402

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 30
dev = torch.device("cuda:0")
tr = SceneTrainer(bench.SPECS, bench.SCENES_TOTAL, dev, seed=0)
gen = torch.Generator(device=dev).manual_seed(2)
ids = torch.arange(64) * 3
batch = bench.make_batch(gen, dev, ids, bench.SCENES_TOTAL)
ids_d = ids.to(dev)
g_rows = 64 * bench.SAMPLES_PER_SCENE
for mode in ("eager", "graph"):
    fn = tr.step if mode == "eager" else tr.step_graphed
    out = {}
    for s in (64, 32, 16, 8):
        b, i = batch[:s * bench.SAMPLES_PER_SCENE], ids_d[:s].contiguous()
        for _ in range(5):
            fn(b, i, epoch=1, global_rows=g_rows)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn(b, i, epoch=1, global_rows=g_rows)
        e1.record()
        torch.cuda.synchronize()
        out[s] = round(e0.elapsed_time(e1) / steps, 4)
    print(mode, "side_stream", tr.side_stream is not None, out, "bound@8", round(out[64] / (8 * out[8]), 3),
          "kernels/step", getattr(tr, "graph_launches", None) if mode == "graph" else "-")
