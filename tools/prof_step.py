#!/usr/bin/env python3
"""Small driver for ncu: a few fused training steps (and one eval forward) of BASELINE config #2 at a configurable batch."""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from deepsdf_b200.trainer import SceneTrainer  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--scenes", type=int, default=64)
ap.add_argument("--steps", type=int, default=2)
ap.add_argument("--latent", type=int, default=256)
args = ap.parse_args()
dev = torch.device("cuda:0")
specs = dict(bench.SPECS)
specs["CodeLength"] = args.latent
tr = SceneTrainer(specs, bench.SCENES_TOTAL, dev, seed=0)
gen = torch.Generator(device=dev).manual_seed(2)
ids = torch.arange(args.scenes) * 3
batch = bench.make_batch(gen, dev, ids, bench.SCENES_TOTAL)
ids_d = ids.to(dev)
for i in range(args.steps):
    loss = tr.step(batch, ids_d, epoch=1)
torch.cuda.synchronize()
# one eval-mode forward over the same points
eng = tr.engine
cb = eng.scene_bias(tr.lat_vecs, ids_d)
ts = tr.tile_scene_for(args.scenes, bench.SAMPLES_PER_SCENE)
xyz = batch[:, :3].contiguous()
sdf = eng.decode_scenes(cb, ts, xyz)
torch.cuda.synchronize()
print("loss", [float(x) for x in loss.cpu()], "sdf mean", float(sdf.mean()))
