#!/usr/bin/env python3
"""Timeline of one tile of the training forward chain (CTA 0, its 4th tile): needs a library built with -DDSDF_TRACE
(DSDF_B200_LIB=ab/libdsdf_trace.so python tools/trace_fwd.py).  Prints SM-clock offsets per layer for epilogue warp 0,
epilogue warp 15 and the MMA warp."""
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from deepsdf_b200 import _lib as L  # noqa: E402
from deepsdf_b200.trainer import SceneTrainer  # noqa: E402

dev = torch.device("cuda:0")
tr = SceneTrainer(bench.SPECS, bench.SCENES_TOTAL, dev, seed=0)
gen = torch.Generator(device=dev).manual_seed(2)
ids = torch.arange(64) * 3
batch = bench.make_batch(gen, dev, ids, bench.SCENES_TOTAL)
ids_d = ids.to(dev)
for _ in range(4):
    tr.step(batch, ids_d, epoch=1)
torch.cuda.synchronize()
which = sys.argv[1] if len(sys.argv) > 1 else "fwd"
fn = getattr(L.lib(), "dsdf_debug_trace" if which == "fwd" else "dsdf_debug_trace_bwd")
buf = np.zeros((6, 16, 40), dtype=np.uint64)
assert fn(buf.ctypes.data_as(C.c_void_p)) == 0
t = buf.astype(np.int64)
fmt = lambda row, ref: " ".join(f"{(int(x) - ref) if x else -1:6d}" for x in row)
for l in range(9):
    if t[0, l, 0] == 0:
        continue
    ref = int(t[0, l, 0])
    print(f"--- layer {l}   (start {ref - int(t[0, 0, 0])})")
    print("epi w0     ", fmt(t[0, l, :16], ref))
    print("epi w15    ", fmt(t[2, l, :16], ref))
    print("epi w0 cta1", fmt(t[4, l, :16], ref))
    print("mma a_ready", fmt(t[1, l, 0:36:2], ref))
    print("mma B full ", fmt(t[1, l, 1:36:2], ref))
    print("tma empty  ", fmt(t[3, l, :18], ref))
print("epilogue events: 0 start, 1 keep words drawn, 2 full[0], q0: 3 computed 4 slot free 5 stored, q1: 6 7 8, 9 full[1], q2: 10 11 12, q3: 13 14 15")
print("mma / tma steps: half 0: aux, K chunks 0..7; half 1: aux, K chunks 0..7  (mma: a_ready satisfied / B stage full; tma: stage empty seen)")
