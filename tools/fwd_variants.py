#!/usr/bin/env python3
"""Forward-kernel time decomposition (config #2 batch): eval / gates only / gates + stash / gates + Philox / full."""
import ctypes as C
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from deepsdf_b200 import _lib as L  # noqa: E402
from deepsdf_b200.trainer import SceneTrainer  # noqa: E402

dev = torch.device("cuda:0")
tr = SceneTrainer(bench.SPECS, bench.SCENES_TOTAL, dev, seed=0)
eng = tr.engine
gen = torch.Generator(device=dev).manual_seed(2)
ids = torch.arange(64) * 3
batch = bench.make_batch(gen, dev, ids, bench.SCENES_TOTAL)
N = batch.shape[0]
b = eng.train_buffers(N, 64)
cb = eng.scene_bias(tr.lat_vecs, ids.to(dev), 1.0)
ts = tr.tile_scene_for(64, bench.SAMPLES_PER_SCENE)
flat = batch.view(-1)


def run(training, dropout_off, acts, gates, reps=10):
    io = L.TrainIO()
    io.b_last = eng._b_last().data_ptr()
    io.training, io.dropout_off = training, dropout_off
    io.seed, io.row0 = 5, 0
    if training:
        io.gt, io.gt_stride, io.clamp, io.inv_n = flat[3:].data_ptr(), 4, 0.1, 1.0 / N
        io.dy_out, io.tile_loss = b["dy"].data_ptr(), b["tile_loss"].data_ptr()
        io.bitmask = b["bitmask"].data_ptr() if gates else None
        io.acts = b["acts"].data_ptr() if acts else None
        io.n_rows_pad = N
    def go():
        L.check(eng.lib.dsdf_fused_forward(C.byref(eng.desc), eng.pack.data_ptr(), cb.data_ptr(), ts.data_ptr(), flat.data_ptr(), 4, N,
                                           b["sdf"].data_ptr(), C.byref(io), eng._st()))
    for _ in range(3):
        go()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(reps):
        go()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


cfgs = [("eval", (0, 0, False, False)), ("train-template, no dropout, no gate store, no stash", (1, 1, False, False)),
        ("gates stored", (1, 1, False, True)), ("gates + stash", (1, 1, True, True)),
        ("Philox, no gate store, no stash", (1, 0, False, False)),
        ("gates + Philox, no stash", (1, 0, False, True)), ("full training forward", (1, 0, True, True))]
res = {}
for rnd in range(4):          # interleaved rounds: the chip warms up and clocks drift under the power cap
    for name, cfg in cfgs:
        res.setdefault(name, []).append(round(run(*cfg), 3))
for name, v in res.items():
    print(f"{name:55s} {v}")
