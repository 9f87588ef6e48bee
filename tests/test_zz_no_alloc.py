"""dsdf_clip_coef_static (csrc/noalloc.cu): the allocation-free variant of dsdf_clip_coef whose 4 KB of scratch is a file-scope
``static __device__`` array inside the library (address resolved once per device with cudaGetSymbolAddress).  No counterpart in
the reference.  Called through the C ABI and compared with the existing entry
point (caller-provided scratch) - the two use the same block ranges and summation order, so they must agree bit for bit - on
two calls in a row (the second finds the first call's partial sums in the shared array), then against torch within the 1e-6
relative tolerance the clip tests use.  Collected last on purpose (the file name).
Runs alone in a fresh process (first use on the device resolves the address) and after the other tests alike."""
import ctypes as C

import pytest
import torch

from deepsdf_b200 import _lib as L

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
MAX_N = 1 << 28           # DSDF_CLIP_STATIC_MAX_N (include/dsdf.h)


def _both(lib, g, max_norm):
    """(existing entry point with caller scratch, static-scratch entry point) -> two [2] tensors (norm, coefficient)."""
    scratch = torch.full((1024,), float("nan"), device=DEV)
    out_ref = torch.zeros(2, device=DEV)
    out_new = torch.zeros(2, device=DEV)
    with torch.cuda.device(DEV):
        st = L.stream_ptr(torch.device(DEV))
        L.check(lib.dsdf_clip_coef(g.data_ptr(), g.numel(), float(max_norm), scratch.data_ptr(), out_ref.data_ptr(), st))
        before = torch.cuda.memory_allocated(DEV), torch.cuda.memory_reserved(DEV)
        L.check(lib.dsdf_clip_coef_static(g.data_ptr(), g.numel(), float(max_norm), out_new.data_ptr(), st))
        assert (torch.cuda.memory_allocated(DEV), torch.cuda.memory_reserved(DEV)) == before      # nothing allocated through torch
    torch.cuda.synchronize()
    return out_ref.cpu(), out_new.cpu()


def test_static_scratch_entry_point_matches_the_workspace_one_on_two_calls_in_a_row():
    lib = L.lib()
    assert hasattr(lib, "dsdf_clip_coef_static")
    gen = torch.Generator(device=DEV).manual_seed(21)
    # call 1: a decoder-gradient-sized vector that gets clipped; call 2: a short one (fewer elements than partial-sum blocks, so
    # most blocks see an empty range and must overwrite the partials call 1 left behind with zeros) that does not
    g1 = torch.randn(1_862_000 + 37, generator=gen, device=DEV) * 0.01
    g2 = torch.randn(700, generator=gen, device=DEV) * 1e-3
    for g, max_norm in ((g1, 1.0), (g2, 1.0), (g1, 50.0)):
        ref, new = _both(lib, g, max_norm)
        assert torch.equal(ref, new), (ref, new)
        nrm = float(torch.linalg.vector_norm(g.double()))
        assert abs(float(new[0]) - nrm) <= 1e-6 * nrm + 1e-12
        coef = min(1.0, max_norm / (nrm + 1e-6))
        assert abs(float(new[1]) - coef) <= 1e-6 * coef
    # the first pair really exercised both branches of the coefficient
    assert float(_both(lib, g1, 1.0)[1][1]) < 1.0 and float(_both(lib, g2, 1.0)[1][1]) == 1.0


def test_static_scratch_entry_point_rejects_what_it_cannot_hold_and_replays_in_a_graph():
    lib = L.lib()
    g = torch.ones(4096, device=DEV)
    out = torch.zeros(2, device=DEV)
    with torch.cuda.device(DEV):
        # beyond the stated maximum: an error with a message, before anything is launched (the pointer is never dereferenced)
        rc = lib.dsdf_clip_coef_static(g.data_ptr(), MAX_N + 1, 1.0, out.data_ptr(), None)
        assert rc != 0 and b"DSDF_CLIP_STATIC_MAX_N" in lib.dsdf_last_error()
        assert lib.dsdf_clip_coef_static(None, 16, 1.0, out.data_ptr(), None) != 0
        assert lib.dsdf_clip_coef_static(g.data_ptr(), 0, 1.0, out.data_ptr(), None) != 0
        # a valid call still works afterwards
        L.check(lib.dsdf_clip_coef_static(g.data_ptr(), g.numel(), 8.0, out.data_ptr(), L.stream_ptr(torch.device(DEV))))
        torch.cuda.synchronize()
        assert abs(float(out[0]) - 64.0) <= 1e-5 and abs(float(out[1]) - 8.0 / (64.0 + 1e-6)) <= 1e-6
        # the use the entry point exists for: inside a captured CUDA graph (nothing to allocate, nothing to hand over)
        out.zero_()
        side = torch.cuda.Stream(device=DEV)
        side.wait_stream(torch.cuda.current_stream(DEV))
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.stream(side):
            with torch.cuda.graph(graph, stream=side):
                L.check(lib.dsdf_clip_coef_static(g.data_ptr(), g.numel(), 8.0, out.data_ptr(), L.stream_ptr(torch.device(DEV))))
        torch.cuda.current_stream(DEV).wait_stream(side)
        g.mul_(2.0)
        graph.replay()
        torch.cuda.synchronize()
        assert abs(float(out[0]) - 128.0) <= 1e-5 and abs(float(out[1]) - 8.0 / (128.0 + 1e-6)) <= 1e-6
        del graph
