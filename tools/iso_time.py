#!/usr/bin/env python3
"""Time the two iso-surface passes on a 512^3 sphere volume (CUDA events, 10 repetitions each)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from deepsdf_b200 import _lib as L  # noqa: E402
from deepsdf_b200.mesh import extract_isosurface  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 512
dev = torch.device("cuda:0")
g = torch.linspace(-1, 1, N, device=dev)
vol = (torch.sqrt(g[:, None, None] ** 2 + g[None, :, None] ** 2 + g[None, None, :] ** 2) - 0.6).contiguous()
lib = L.lib()
nb = int(lib.dsdf_iso_num_blocks(N, N, N))
counts = torch.empty(nb, dtype=torch.int32, device=dev)
st = L.stream_ptr(dev)


def timed(f, reps=10):
    f()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        f()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


t_count = timed(lambda: L.check(lib.dsdf_iso_count(vol.data_ptr(), N, N, N, 0.0, counts.data_ptr(), st)))
t_all = timed(lambda: extract_isosurface(vol, (-1, -1, -1), 2.0 / (N - 1)))
gb = vol.numel() * 4 / 1e9
print(f"N={N}: count pass {t_count:.3f} ms ({gb / t_count * 1e3:.0f} GB/s of the volume), whole extraction {t_all:.3f} ms, "
      f"{extract_isosurface(vol, (-1, -1, -1), 2.0 / (N - 1)).shape[0]} triangles")
