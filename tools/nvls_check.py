#!/usr/bin/env python3
"""dsdf_nvls_allreduce_f32 against ncclAllReduce under torchrun: same sums (to fp32 summation order) on random gradients of the
flat decoder buffer's size, repeated back to back (barrier channels must come back to zero), and the time of each."""
import os
import sys

import torch
import torch.distributed as dist
import torch.distributed._symmetric_memory as symm

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from deepsdf_b200 import _lib as L  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
lib = L.lib()
n = 1843264
buf = symm.empty(n, dtype=torch.float32, device=dev)
buf.zero_()
hdl = symm.rendezvous(buf, dist.group.WORLD.group_name)
assert hdl.multicast_ptr != 0 and hdl.signal_pad_size >= 4 * lib.dsdf_nvls_signal_words(world), (hdl.multicast_ptr, hdl.signal_pad_size)
gen = torch.Generator(device=dev).manual_seed(100 + rank)
worst = 0.0
for it in range(6):
    x = torch.randn(n, device=dev, generator=gen) * (1.0 + it)
    ref = x.clone()
    dist.all_reduce(ref)
    buf.copy_(x)
    with torch.cuda.device(dev):
        L.check(lib.dsdf_nvls_allreduce_f32(hdl.multicast_ptr, hdl.signal_pad_ptrs_dev, hdl.rank, world, n, L.stream_ptr(dev)))
    torch.cuda.synchronize()
    worst = max(worst, float((buf - ref).abs().max() / ref.abs().max()))


def timeit(fn, reps=100):
    for _ in range(5):
        fn()
    dist.barrier(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    t = torch.tensor([a.elapsed_time(b) / reps], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t) * 1e3


buf.zero_()
y = torch.zeros(n, device=dev)
st = L.stream_ptr(dev)
t_own = timeit(lambda: L.check(lib.dsdf_nvls_allreduce_f32(hdl.multicast_ptr, hdl.signal_pad_ptrs_dev, hdl.rank, world, n, st)))
t_nccl = timeit(lambda: dist.all_reduce(y))
pad = hdl.get_signal_pad(hdl.rank, (lib.dsdf_nvls_signal_words(world),), torch.int32)
w = torch.tensor([worst], device=dev); dist.all_reduce(w, op=dist.ReduceOp.MAX)
if rank == 0:
    print(f"{world} ranks, {n * 4 / 1e6:.1f} MB: own NVLS all-reduce {t_own:.1f} us, ncclAllReduce {t_nccl:.1f} us; "
          f"max rel diff vs NCCL {float(w):.2e}; signal pad ({hdl.signal_pad_size} B) back to zero: {bool((pad == 0).all())}")
torch.cuda.synchronize()
dist.barrier()
dist.destroy_process_group()
