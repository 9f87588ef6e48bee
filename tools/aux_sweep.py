"""Bring-up diagnostic: which (LBO, SBO) make the SWIZZLE_NONE K-major aux step of csrc/selftest.cu agree with a matmul."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from deepsdf_b200 import _lib as L

lib, dev = L.lib(), "cuda:0"
g = torch.Generator().manual_seed(2)
K = 128
A = (torch.randn(256, K, generator=g) * 0.5).to(torch.bfloat16)
B = (torch.randn(256, K, generator=g) * 0.5).to(torch.bfloat16)
Ax = (torch.randn(256, 32, generator=g) * 0.5).to(torch.bfloat16)
Bx = (torch.randn(256, 32, generator=g) * 0.5).to(torch.bfloat16)
Ad, Bd, Axd, Bxd = A.to(dev), B.to(dev), Ax.to(dev), Bx.to(dev)
for aux_k in (16, 32):
    ref = A.float() @ B.float().t() + Ax[:, :aux_k].float() @ Bx[:, :aux_k].float().t()
    for lbo, sbo in ((2048, 128), (128, 2048), (2048, 256), (256, 2048), (4096, 128), (16, 128)):
        C = torch.zeros(256, 256, device=dev)
        L.check(lib.dsdf_selftest_umma_aux(Ad.data_ptr(), Bd.data_ptr(), Axd.data_ptr(), Bxd.data_ptr(), C.data_ptr(), K, aux_k, lbo, sbo, None))
        torch.cuda.synchronize()
        print(f"aux_k {aux_k} lbo {lbo} sbo {sbo}: max err {(C.cpu() - ref).abs().max().item():.4f}", flush=True)
