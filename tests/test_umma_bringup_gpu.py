"""Descriptor conventions the fused kernels rely on, checked against a PyTorch matmul on one CTA pair (csrc/selftest.cu):
the cta_group::2 SW128 GEMM, and the aux K = 16 / 32 step in the SWIZZLE_NONE K-major canonical layout (bias / xyz / short
latent columns of the forward chain: thread-written A block, no-swizzle TMA boxes for B, LBO = 2048, SBO = 128)."""
import pytest
import torch

from deepsdf_b200 import _lib as L

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _operands(K, seed):
    g = torch.Generator().manual_seed(seed)
    A = (torch.randn(256, K, generator=g) * 0.5).to(torch.bfloat16)
    B = (torch.randn(256, K, generator=g) * 0.5).to(torch.bfloat16)
    Ax = (torch.randn(256, 32, generator=g) * 0.5).to(torch.bfloat16)
    Bx = (torch.randn(256, 32, generator=g) * 0.5).to(torch.bfloat16)
    return A, B, Ax, Bx


@pytest.mark.parametrize("K", [64, 256])
def test_pair_gemm_sw128(K):
    lib = L.lib()
    A, B, _, _ = _operands(K, 1)
    C = torch.zeros(256, 256, device=DEV)
    Ad, Bd = A.to(DEV), B.to(DEV)
    L.check(lib.dsdf_selftest_umma_pair(0, Ad.data_ptr(), Bd.data_ptr(), C.data_ptr(), K, None))
    torch.cuda.synchronize()
    ref = A.float() @ B.float().t()
    assert (C.cpu() - ref).abs().max().item() <= 1e-3 * ref.abs().max().item()


@pytest.mark.parametrize("aux_k", [16, 32])
def test_aux_step_noswizzle_kmajor(aux_k):
    lib = L.lib()
    K = 128
    A, B, Ax, Bx = _operands(K, 2)
    Ad, Bd, Axd, Bxd = A.to(DEV), B.to(DEV), Ax.to(DEV), Bx.to(DEV)
    ref = A.float() @ B.float().t() + Ax[:, :aux_k].float() @ Bx[:, :aux_k].float().t()
    C = torch.zeros(256, 256, device=DEV)
    L.check(lib.dsdf_selftest_umma_aux(Ad.data_ptr(), Bd.data_ptr(), Axd.data_ptr(), Bxd.data_ptr(), C.data_ptr(), K, aux_k, 2048, 128, None))
    torch.cuda.synchronize()
    err = (C.cpu() - ref).abs().max().item()
    print(f"aux_k {aux_k}: LBO 2048 / SBO 128 -> max err {err:.3e} (|ref| max {ref.abs().max().item():.2f})")
    assert err <= 1e-3 * ref.abs().max().item()
