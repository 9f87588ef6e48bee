"""Grid evaluation of ``deep_sdf/mesh.py:26-73`` ``create_mesh`` (the SDF volume), the iso-surface + ply tail of it
(``convert_sdf_samples_to_ply``, mesh.py:88-155, on the device) and the optional derivative outputs the optimisation / analysis consumers need
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
                    slab: Optional[Tuple[int, int]] = None, out_device=None, latent_tangents: Optional[torch.Tensor] = None,
                    exact: bool = False):
    """SDF values of create_mesh's grid, shape [n_slab, N, N] (``slab=None``: the whole [N, N, N] volume).

    exact=True (8x512-class decoders, SDF only): the split-bf16 tensor-core mode, ~1e-6 against the reference's fp32
    arithmetic instead of the bf16 chain's ~1e-3, at about 4x the time (small decoders are fp32 anyway).

    with_grad=True additionally returns (dxyz [n, N, N, 3], dlat [n, N, N, L]); with ``latent_tangents`` [L, T] the
    latent part is contracted on the fly to [n, N, N, T] (what ``jvp`` with T tangents would give) so that the
    L-wide gradient never has to be stored.  Results stay on the device unless ``out_device`` says otherwise."""
    eng = fused_engine_for(decoder)
    if eng is None:
        from .small import small_engine_for
        sm = small_engine_for(decoder)
        if sm is None:
            raise L.DsdfError("create_mesh_sdf needs a decoder of the fused 8x512 class or of the small-network class (hidden width "
                              "<= 128); evaluate other NetworkSpecs through Decoder.forward chunks")
        return _create_mesh_sdf_small(sm, latent_vec, N, max_batch, with_grad, slab, out_device, latent_tangents)
    dec_ = decoder.module if isinstance(decoder, torch.nn.DataParallel) else decoder
    exact = bool(exact or (getattr(dec_, "fused_exact", False) and not with_grad))
    if exact and with_grad:
        raise L.DsdfError("create_mesh_sdf(exact=True) evaluates the SDF only; take derivatives from the bf16 path (with_grad=True)")
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
    zero_id = torch.zeros(1, dtype=torch.int64, device=dev)
    cb = eng.scene_bias_exact(table, zero_id) if exact else eng.scene_bias(table, zero_id)
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
            eng.decode_scenes(cb, ts[:n_pad // TILE], xyz[:n], out=sdf_c, exact=exact)
        sdf[head:head + n].copy_(sdf_c[:n])
        head += n
    shape = (i1 - i0, N, N)
    if not with_grad:
        return sdf.view(shape)
    return sdf.view(shape), dxyz.view(*shape, 3), dlat.view(*shape, -1)


def _create_mesh_sdf_small(sm, latent_vec, N, max_batch, with_grad, slab, out_device, latent_tangents):
    """create_mesh_sdf for the small decoders: the same chunk loop on the fused fp32 small-network kernel (one launch per
    chunk evaluates SDF and, with derivatives, G = dSDF/d[latent, xyz])."""
    import ctypes as C
    dev, Lz, D = sm.device, sm.spec.latent_size, sm.spec.geom_dimension
    if D != 3:
        raise L.DsdfError("create_mesh_sdf evaluates a 3-D grid (geom_dimension 3)")
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
    chunk = max(TILE, min(int(max_batch), total))
    xyz = torch.empty(chunk, 3, device=dev)
    sdf_c = torch.empty(chunk, device=dev)
    G = torch.empty(chunk, Lz + 3, device=dev) if with_grad else None
    lib = L.lib()
    head = 0
    while head < total:
        n = min(chunk, total - head)
        with torch.cuda.device(dev):
            L.check(lib.dsdf_grid_points(N, first + head, n, xyz.data_ptr(), L.stream_ptr(dev)))
        sm.sdf_and_input_grad(xyz[:n], latent=latent_vec, with_grad=with_grad, sdf=sdf_c, G=G)
        if with_grad:
            dxyz[head:head + n].copy_(G[:n, Lz:Lz + 3])
            dlat[head:head + n].copy_(G[:n, :Lz] if latent_tangents is None else G[:n, :Lz] @ latent_tangents)
        sdf[head:head + n].copy_(sdf_c[:n])
        head += n
    shape = (i1 - i0, N, N)
    if not with_grad:
        return sdf.view(shape)
    return sdf.view(shape), dxyz.view(*shape, 3), dlat.view(*shape, -1)


# ---------------------------------------------------------------------------------------------- iso-surface + ply
def extract_isosurface(volume: torch.Tensor, voxel_grid_origin=(-1.0, -1.0, -1.0), voxel_size=1.0, level: float = 0.0) -> torch.Tensor:
    """Triangle soup [T, 3, 3] (device) of the ``level`` set of ``volume`` [nx, ny, nz] (fp32 CUDA tensor in create_mesh's
    ordering): vertex = voxel_grid_origin + grid_coordinate * voxel_size, normals towards increasing SDF.  Marching
    tetrahedra in ``libdsdf_b200.so`` (``csrc/isosurface.cu``) - the role of ``skimage.measure.marching_cubes`` in
    ``convert_sdf_samples_to_ply`` (deep_sdf/mesh.py:113-115), with the volume never leaving the device."""
    import ctypes as C
    L.require_cuda(volume)
    if volume.dim() != 3 or volume.dtype != torch.float32:
        raise L.DsdfError("extract_isosurface needs a 3-D fp32 volume")
    volume = volume.contiguous()
    nx, ny, nz = (int(s) for s in volume.shape)
    if not isinstance(voxel_size, (list, tuple)):
        voxel_size = [voxel_size] * 3                                            # mesh.py:111-112
    lib, dev = L.lib(), volume.device
    nb = int(lib.dsdf_iso_num_blocks(nx, ny, nz))
    if nb == 0:
        return torch.zeros(0, 3, 3, device=dev)
    counts = torch.empty(nb, dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        L.check(lib.dsdf_iso_count(volume.data_ptr(), nx, ny, nz, float(level), counts.data_ptr(), L.stream_ptr(dev)))
        incl = torch.cumsum(counts, 0, dtype=torch.int64)                        # plumbing: scan of the per-block counts
        offsets = (incl - counts).contiguous()
        total = int(incl[-1])
        verts = torch.empty(total, 3, 3, device=dev)
        if total:
            o3 = (C.c_float * 3)(*[float(x) for x in voxel_grid_origin])
            v3 = (C.c_float * 3)(*[float(x) for x in voxel_size])
            L.check(lib.dsdf_iso_emit(volume.data_ptr(), nx, ny, nz, float(level), o3, v3, offsets.data_ptr(), verts.data_ptr(),
                                      L.stream_ptr(dev)))
    return verts


def write_ply(ply_filename_out: str, triangles, offset=None, scale=None) -> None:
    """Binary little-endian ply with the elements ``convert_sdf_samples_to_ply`` writes through plyfile (mesh.py:131-146):
    ``vertex`` (float x, y, z) and ``face`` (list uchar int vertex_indices); ``scale`` / ``offset`` as in mesh.py:123-127."""
    import numpy as np
    tri = triangles.detach().cpu().numpy() if isinstance(triangles, torch.Tensor) else np.asarray(triangles)
    pts = tri.reshape(-1, 3).astype(np.float32)
    if scale is not None:
        pts = pts / scale
    if offset is not None:
        pts = pts - offset
    n_faces = pts.shape[0] // 3
    verts = np.zeros(pts.shape[0], dtype=[("x", "<f4"), ("y", "<f4"), ("z", "<f4")])
    verts["x"], verts["y"], verts["z"] = pts[:, 0], pts[:, 1], pts[:, 2]
    faces = np.zeros(n_faces, dtype=[("n", "u1"), ("vertex_indices", "<i4", (3,))])
    faces["n"] = 3
    faces["vertex_indices"] = np.arange(3 * n_faces, dtype=np.int32).reshape(n_faces, 3)
    header = ("ply\nformat binary_little_endian 1.0\nelement vertex {}\nproperty float x\nproperty float y\nproperty float z\n"
              "element face {}\nproperty list uchar int vertex_indices\nend_header\n").format(pts.shape[0], n_faces)
    with open(ply_filename_out, "wb") as f:
        f.write(header.encode("ascii"))
        f.write(verts.tobytes())
        f.write(faces.tobytes())


def convert_sdf_samples_to_ply(pytorch_3d_sdf_tensor: torch.Tensor, voxel_grid_origin, voxel_size, ply_filename_out: str,
                               offset=None, scale=None) -> int:
    """Same signature as deep_sdf/mesh.py:88-95, for a DEVICE tensor (create_mesh_sdf's result): iso-surface at level 0 on
    the GPU, then the ply file.  Returns the number of triangles."""
    tri = extract_isosurface(pytorch_3d_sdf_tensor.to(torch.float32), voxel_grid_origin, voxel_size, 0.0)
    write_ply(ply_filename_out, tri, offset, scale)
    return int(tri.shape[0])


def create_mesh(decoder, latent_vec: torch.Tensor, filename: str, N: int = 256, max_batch: int = 1 << 21, offset=None, scale=None) -> int:
    """deep_sdf/mesh.py:26-85 end to end on the device: grid evaluation (fused forward) + iso-surface + ply file.  Like this
    fork's reference the mesh is written to ``filename`` AS GIVEN (mesh.py:31,76-83: ``ply_filename = filename``; callers such
    as ``create_mesh_from_latent`` pass a name that already ends in .ply) and the decoder is left in eval mode (mesh.py:29)."""
    decoder.eval()
    sdf = create_mesh_sdf(decoder, latent_vec, N=N, max_batch=max_batch)
    voxel_size = 2.0 / (N - 1)                                                   # mesh.py:33-35
    return convert_sdf_samples_to_ply(sdf, [-1.0, -1.0, -1.0], voxel_size, filename, offset, scale)


# ---------------------------------------------------------------------------------------------- microstructure evaluation
location_lookup = {"x0": (0, -1), "x1": (0, 1), "y0": (1, -1), "y1": (1, 1), "z0": (2, -1), "z1": (2, 1)}   # mesh.py:157-164


def evaluate_network_sdf(decoder, lat_vec_red: torch.Tensor, samples_orig: torch.Tensor, tiling, cap_border_dict=None,
                         with_grad: bool = False):
    """The decoder part of ``prepare_samples`` + ``evaluate_network`` (deep_sdf/mesh.py:457-514) for per-vertex latent codes:

      samples   = tiling transform of ``samples_orig`` [N_tot, >=3] into one unit cell              (mesh.py:470-477)
      sdf       = decode_sdf(decoder, None, hstack([lat_vec_red, samples]))                          (mesh.py:490-495)
      sdf       = border caps of ``cap_border_dict`` ({"x0": {"cap": +-1, "measure": m}, ...}) and the unit-cube cap  (:499-514)

    Returns ``(sdf [N_tot], dsdf_dlat [N_tot, L] or None)``.  ``dsdf_dlat`` is what the per-(latent, control point) jvp loop
    of mesh.py:405-422 needs: d sdf / d control point c of code component i = ``dsdf_dlat[:, i] * basis[:, c]`` (zero where a
    cap decides the value).  FlexiCubes / splinepy (the mesh from ``sdf`` and the spline basis) are not part of this path.
    Three kernels of libdsdf_b200.so around the fused forward / dgrad / input-gradient launches; nothing leaves the device."""
    import ctypes as C
    from .small import small_engine_for
    sm = small_engine_for(decoder)                       # 4x32 / 4x64 / 6x128: the networks the optimisation runs use
    eng = None if sm is not None else fused_engine_for(decoder)
    import ctypes as _C
    if sm is None and (eng is None or L.lib().dsdf_fused_rows_supported(_C.byref(eng.desc)) != 1):
        raise L.DsdfError("evaluate_network_sdf needs a small-network-class decoder (hidden width <= 128) or an 8x512-class decoder "
                          "whose per-row codes the fused row mode supports (CodeLength <= 16, or 256)")
    if (sm or eng).spec.geom_dimension != 3:
        raise L.DsdfError("evaluate_network_sdf works on 3-D samples (geom_dimension 3)")
    dev, Lz = (sm or eng).device, (sm or eng).spec.latent_size
    so = samples_orig.detach().to(dev, torch.float32).contiguous()
    lat = lat_vec_red.detach().to(dev, torch.float32).contiguous()
    n = so.shape[0]
    if lat.shape != (n, Lz) or so.shape[1] < 3:
        raise L.DsdfError("lat_vec_red must be [N_tot, L] and samples_orig [N_tot, >= 3]")
    n_pad = (n + TILE - 1) // TILE * TILE
    queries = torch.empty(n_pad, Lz + 3, device=dev)
    lib, st = L.lib(), L.stream_ptr(dev)
    t3 = (C.c_float * 3)(*[float(t) for t in tiling])
    with torch.cuda.device(dev):
        L.check(lib.dsdf_tile_queries(so.data_ptr(), so.stride(0), lat.data_ptr(), lat.stride(0), Lz, n, n_pad, t3, queries.data_ptr(), st))
    if sm is not None:
        sdf, G = sm.sdf_and_input_grad(queries[:n], with_grad=with_grad)
    else:
        sdf, G = eng.sdf_and_input_grad_rows(queries, with_grad=with_grad)
    caps = list((cap_border_dict or {}).items())
    if len(caps) > 6 or any(k not in location_lookup for k, _ in caps):
        raise L.DsdfError("cap_border_dict takes the six keys of location_lookup (x0, x1, y0, y1, z0, z1)")
    dims = (C.c_int32 * 6)(*([location_lookup[k][0] for k, _ in caps] + [0] * (6 - len(caps))))
    mult = (C.c_double * 6)(*([float(location_lookup[k][1]) for k, _ in caps] + [0.0] * (6 - len(caps))))
    meas = (C.c_double * 6)(*([float(v["measure"]) for _, v in caps] + [0.0] * (6 - len(caps))))
    capv = (C.c_int32 * 6)(*([int(v["cap"]) for _, v in caps] + [1] * (6 - len(caps))))
    for _, v in caps:
        if v["cap"] not in (-1, 1):
            raise ValueError("Cap must be -1 or 1")                                   # mesh.py:509
    with torch.cuda.device(dev):
        L.check(lib.dsdf_cap_sdf(sdf.data_ptr(), G.data_ptr() if G is not None else None, G.stride(0) if G is not None else 0, Lz,
                                 so.data_ptr(), so.stride(0), n, len(caps), dims, mult, meas, capv, st))
    return sdf[:n], (G[:n, :Lz] if G is not None else None)
