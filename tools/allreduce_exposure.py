#!/usr/bin/env python3
"""Where the gradient all-reduce sits in the strong-scaled step (fixed 64-scene batch split over the ranks): under torchrun,
every rank times, with CUDA events on its compute stream, (a) the whole step and (b) the span from the point the all-reduce is
issued (finish has produced the flat gradient) to the point the stream has joined it (latent Adam, which needs no exchange,
runs inside that span).  Max over ranks, eager launches.
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 tools/allreduce_exposure.py"""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from deepsdf_b200.trainer import SceneTrainer  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
tr = SceneTrainer(bench.SPECS, bench.SCENES_TOTAL, dev, rank=rank, world_size=world, seed=0)
s_loc = bench.SCENES_PER_STEP // world
gen = torch.Generator(device=dev).manual_seed(2 + rank)
ids = (torch.arange(s_loc) * 3).sort().values
batch = bench.make_batch(gen, dev, ids, bench.SCENES_TOTAL)
ids_d = ids.to(dev)
g_rows = bench.SCENES_PER_STEP * bench.SAMPLES_PER_SCENE
for _ in range(5):
    tr.step(batch, ids_d, epoch=1, global_rows=g_rows, row0=rank * s_loc * bench.SAMPLES_PER_SCENE)
dist.barrier(); torch.cuda.synchronize()
steps = 30
tr.comm_events = []
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(steps):
    tr.step(batch, ids_d, epoch=1, global_rows=g_rows, row0=rank * s_loc * bench.SAMPLES_PER_SCENE)
b.record()
torch.cuda.synchronize()
span = sum(x.elapsed_time(y) for x, y in tr.comm_events) / len(tr.comm_events)
t = torch.tensor([a.elapsed_time(b) / steps, span], device=dev, dtype=torch.float64)
dist.all_reduce(t, op=dist.ReduceOp.MAX)
# the same span without a collective: one rank alone (latent Adam only)
tr1 = SceneTrainer(bench.SPECS, bench.SCENES_TOTAL, dev, rank=0, world_size=1, seed=0)
for _ in range(3):
    tr1.step(batch, ids_d, epoch=1, global_rows=g_rows)
tr1.comm_events = []
torch.cuda.synchronize()
a.record()
for _ in range(steps):
    tr1.step(batch, ids_d, epoch=1, global_rows=g_rows)
b.record()
torch.cuda.synchronize()
span1 = sum(x.elapsed_time(y) for x, y in tr1.comm_events) / len(tr1.comm_events)
t1 = torch.tensor([a.elapsed_time(b) / steps, span1], device=dev, dtype=torch.float64)
dist.all_reduce(t1, op=dist.ReduceOp.MAX)
if rank == 0:
    print(f"{world} ranks, {s_loc} scenes each: step {t[0]:.4f} ms, all-reduce issue -> join {t[1] * 1e3:.1f} us; "
          f"the same shard without the collective: step {t1[0]:.4f} ms, span {t1[1] * 1e3:.1f} us "
          f"=> exposed all-reduce ~{(t[1] - t1[1]) * 1e3:.1f} us of the {t[0] * 1e3:.0f} us step")
torch.cuda.synchronize()
dist.barrier()
dist.destroy_process_group()
