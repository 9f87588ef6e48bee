#!/usr/bin/env python3
"""GPU bring-up of the CTA-pair (cta_group::2) conventions: dsdf_selftest_umma_pair vs torch.matmul."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from deepsdf_b200 import _lib as L  # noqa: E402

dev = torch.device("cuda:0")
lib = L.lib()
L.check(lib.dsdf_check_device())
ok_all = True
for mode in (0, 4):
    for K in (64, 128, 256):
        g = torch.Generator().manual_seed(K + mode)
        A = torch.randn(256, K, generator=g).to(torch.bfloat16)
        B = torch.randn(256, K, generator=g).to(torch.bfloat16)
        ref = A.float() @ B.float().t()
        Ad, Bd = A.to(dev), B.to(dev)
        C = torch.full((256, 256), float("nan"), device=dev)
        rc = lib.dsdf_selftest_umma_pair(mode, Ad.data_ptr(), Bd.data_ptr(), C.data_ptr(), K, L.stream_ptr(dev))
        if rc != 0:
            print("launch error", lib.dsdf_last_error().decode()); ok_all = False; continue
        try:
            torch.cuda.synchronize()
        except Exception as e:  # noqa: BLE001
            print(f"[pair mode={mode} K={K}] CUDA error: {e}"); ok_all = False; break
        Cc = C.cpu()
        err = (Cc - ref).abs()
        q = [float(err[r0:r0 + 128, c0:c0 + 128].max()) for r0 in (0, 128) for c0 in (0, 128)]
        ok = float(err.max()) < 1e-2 * float(ref.abs().max())
        ok_all &= ok
        print(f"[pair mode={mode} K={K}] max_err={float(err.max()):.3g} per-quadrant(rows x cols)={['%.3g' % x for x in q]} nan={int(torch.isnan(Cc).sum())} -> {'OK' if ok else 'MISMATCH'}")
print("ALL OK" if ok_all else "FAILED")
