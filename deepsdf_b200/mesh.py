"""Grid evaluation of ``deep_sdf/mesh.py:26-73`` ``create_mesh`` (the SDF volume; marching cubes / ply export stay on
the host and are out of scope) with the optional derivative outputs the optimisation / analysis consumers need
(``deep_sdf/mesh.py:405-454``): dSDF/dxyz and dSDF/dlatent for every grid point.

The N^3 grid is generated on the device chunk by chunk (``dsdf_grid_points``), every chunk runs the fused forward (and,
with derivatives, the dgrad chain + the input-gradient GEMM); nothing crosses PCIe but the result the caller asks for.
Multi-GPU: ``slab=(i0, i1)`` restricts the evaluation to the contiguous range of the slowest grid index (z-slabs of
the reference's ordering, ``mesh.py:45-47``) - ranks evaluate disjoint slabs with no collective.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from . import _lib as L
from .sharding import slab_for_rank  # noqa: F401  (re-exported: callers shard the grid with it)
from .utils import fused_engine_for

TILE = 128


def create_mesh_sdf(decoder, latent_vec: torch.Tensor, N: int = 256, max_batch: int = 1 << 21, with_grad: bool = False,
                    slab: Optional[Tuple[int, int]] = None, out_device=None, latent_tangents: Optional[torch.Tensor] = None):
    """SDF values of create_mesh's grid, shape [n_slab, N, N] (``slab=None``: the whole [N, N, N] volume).

    with_grad=True additionally returns (dxyz [n, N, N, 3], dlat [n, N, N, L]); with ``latent_tangents`` [L, T] the
    latent part is contracted on the fly to [n, N, N, T] (what ``jvp`` with T tangents would give) so that the
    L-wide gradient never has to be stored.  Results stay on the device unless ``out_device`` says otherwise."""
    eng = fused_engine_for(decoder)
    if eng is None:
        raise L.DsdfError("create_mesh_sdf needs a decoder supported by the fused path (8x512 class); use Decoder.forward "
                          "chunks for the small networks")
    dev = eng.device
    Lz = eng.spec.latent_size
    i0, i1 = slab if slab is not None else (0, N)
    first, total = i0 * N * N, (i1 - i0) * N * N
    out_device = out_device or dev
    sdf = torch.empty(total, device=out_device)
    dxyz = dlat = None
    if with_grad:
        T = Lz if latent_tangents is None else latent_tangents.shape[1]
        dxyz = torch.empty(total, 3, device=out_device)
        dlat = torch.empty(total, T, device=out_device)
        if latent_tangents is not None:
            latent_tangents = latent_tangents.to(dev, torch.float32)
    table = latent_vec.detach().reshape(1, -1).to(dev, torch.float32).contiguous()
    cb = eng.scene_bias(table, torch.zeros(1, dtype=torch.int64, device=dev))
    chunk = max(TILE, (int(max_batch) // TILE) * TILE)
    chunk = min(chunk, (total + TILE - 1) // TILE * TILE)
    xyz = torch.zeros(chunk, 3, device=dev)
    ts = torch.zeros(chunk // TILE, dtype=torch.int32, device=dev)
    sdf_c = torch.empty(chunk, device=dev)
    G = torch.empty(chunk, Lz + 3, device=dev) if with_grad else None
    head = 0
    while head < total:
        n = min(chunk, total - head)
        eng.grid_points(N, first + head, n, out=xyz)
        if with_grad:
            eng.sdf_and_input_grad(cb, ts, xyz, G=G, sdf=sdf_c)           # rows >= n hold stale points: computed, not copied
            dxyz[head:head + n].copy_(G[:n, Lz:Lz + 3])
            if latent_tangents is None:
                dlat[head:head + n].copy_(G[:n, :Lz])
            else:
                torch.matmul(G[:n, :Lz], latent_tangents, out=dlat[head:head + n]) if dlat.device == dev else \
                    dlat[head:head + n].copy_(G[:n, :Lz] @ latent_tangents)
        else:
            n_pad = (n + TILE - 1) // TILE * TILE
            eng.decode_scenes(cb, ts[:n_pad // TILE], xyz[:n], out=sdf_c)
        sdf[head:head + n].copy_(sdf_c[:n])
        head += n
    shape = (i1 - i0, N, N)
    if not with_grad:
        return sdf.view(shape)
    return sdf.view(shape), dxyz.view(*shape, 3), dlat.view(*shape, -1)
