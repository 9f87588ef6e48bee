#!/usr/bin/env python3
"""tcgen05.mma issue-to-completion rate for the operand layouts the kernels use (single CTA, M128 x N x K16 steps)."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from deepsdf_b200 import _lib as L
dev = torch.device("cuda:0"); lib = L.lib()
reps = 64
for mode, tag in ((0, "K-major A,B"), (3, "MN-major A,B"), (1, "A MN-major"), (2, "B MN-major")):
    for K, N in ((256, 256), (128, 256), (256, 128)):
        a_mn, b_mn = mode & 1, mode & 2
        A = torch.randn(128, K).to(torch.bfloat16); B = torch.randn(N, K).to(torch.bfloat16)
        Ad = (A.t().contiguous() if a_mn else A).to(dev); Bd = (B.t().contiguous() if b_mn else B).to(dev)
        C = torch.zeros(128, N, device=dev); cyc = torch.zeros(1, dtype=torch.int64, device=dev)
        al = K * 128 if a_mn else 16; bl = K * 128 if b_mn else 16
        rc = lib.dsdf_selftest_umma(mode | (reps << 8), Ad.data_ptr(), Bd.data_ptr(), C.data_ptr(), cyc.data_ptr(), K, N, al, 1024, bl, 1024, L.stream_ptr(dev))
        torch.cuda.synchronize()
        n_mma = reps * K // 16
        print(f"{tag:14s} K={K} N={N}: {int(cyc)} cycles for {n_mma} MMAs -> {int(cyc) / n_mma:.1f} cycles/MMA (nominal {128 * N // 256})  rc={rc}")
