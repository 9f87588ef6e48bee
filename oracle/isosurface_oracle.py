"""TEST INFRASTRUCTURE ONLY (imported by tests/, never by the product): CPU restatement of ``csrc/isosurface.cu``.

The reference's iso-surface step is ``skimage.measure.marching_cubes`` (deep_sdf/mesh.py:113-115; scikit-image is a
third-party dependency the reference does not pin, absent in this image), i.e. Lewiner's marching cubes.  The device
kernel is marching TETRAHEDRA (six Kuhn tetrahedra per cube), a different triangulation of the same level set, so this
oracle pins the KERNEL (bit-exact triangle soup, fp32 arithmetic step by step), and the tests pin the GEOMETRY against
analytic shapes (vertices on the level set, area, enclosed volume, watertightness).  PARITY WITH skimage IS UNPINNED.
"""
import numpy as np

TETS = [(0, 1, 3, 7), (0, 1, 5, 7), (0, 2, 3, 7), (0, 2, 6, 7), (0, 4, 5, 7), (0, 4, 6, 7)]
f32 = np.float32


def _corner(b):
    return (b & 1, (b >> 1) & 1, b >> 2)


def _edge_vertex(v, ijk, ca, cb, level):
    a, b = _corner(ca), _corner(cb)
    lo, hi = (ca, cb) if a < b else (cb, ca)                     # lexicographic (x, y, z) = order of the linear grid index
    vl, vh = f32(v[lo]), f32(v[hi])
    t = f32(f32(level) - vl) / f32(vh - vl)
    l, h = _corner(lo), _corner(hi)
    return [f32(f32(ijk[k] + l[k]) + (t if h[k] != l[k] else f32(0.0))) for k in range(3)]


def _triangle(p0, p1, p2, r, origin, voxel):
    u = [f32(p1[k] - p0[k]) for k in range(3)]
    w = [f32(p2[k] - p0[k]) for k in range(3)]
    n = [f32(f32(u[1] * w[2]) - f32(u[2] * w[1])), f32(f32(u[2] * w[0]) - f32(u[0] * w[2])), f32(f32(u[0] * w[1]) - f32(u[1] * w[0]))]
    d = f32(f32(f32(n[0] * f32(r[0])) + f32(n[1] * f32(r[1]))) + f32(n[2] * f32(r[2])))
    if d < 0:
        p1, p2 = p2, p1
    return [[f32(f32(origin[k]) + f32(p[k] * f32(voxel[k]))) for k in range(3)] for p in (p0, p1, p2)]


def marching_tetrahedra(volume, origin=(-1.0, -1.0, -1.0), voxel=(1.0, 1.0, 1.0), level=0.0):
    """Triangle soup [T, 3, 3] fp32 in the kernel's order: cubes by linear index (x slowest), tetrahedra 0..5, triangles."""
    vol = np.asarray(volume, dtype=np.float32)
    nx, ny, nz = vol.shape
    out = []
    for ix in range(nx - 1):
        for iy in range(ny - 1):
            for iz in range(nz - 1):
                v = [vol[ix + (b & 1), iy + ((b >> 1) & 1), iz + (b >> 2)] for b in range(8)]
                inside = [x < f32(level) for x in v]
                if all(inside) or not any(inside):
                    continue
                ijk = (ix, iy, iz)
                for tet in TETS:
                    ins = [c for c in tet if inside[c]]
                    outs = [c for c in tet if not inside[c]]
                    if len(ins) in (0, 4):
                        continue
                    direction = lambda a, b: [_corner(b)[k] - _corner(a)[k] for k in range(3)]
                    if len(ins) == 1 or len(ins) == 3:
                        lone = ins[0] if len(ins) == 1 else outs[0]
                        rest = [c for c in tet if c != lone]
                        r = direction(lone, rest[0]) if len(ins) == 1 else direction(rest[0], lone)
                        ps = [_edge_vertex(v, ijk, lone, c, level) for c in rest]
                        out.append(_triangle(ps[0], ps[1], ps[2], r, origin, voxel))
                    else:
                        r = direction(ins[0], outs[0])
                        q0 = _edge_vertex(v, ijk, ins[0], outs[0], level)
                        q1 = _edge_vertex(v, ijk, ins[0], outs[1], level)
                        q2 = _edge_vertex(v, ijk, ins[1], outs[1], level)
                        q3 = _edge_vertex(v, ijk, ins[1], outs[0], level)
                        out.append(_triangle(q0, q1, q2, r, origin, voxel))
                        out.append(_triangle(q0, q2, q3, r, origin, voxel))
    return np.asarray(out, dtype=np.float32).reshape(-1, 3, 3)


def mesh_area_volume(tri):
    """Area and (divergence theorem) enclosed volume of an oriented triangle soup [T,3,3], float64."""
    t = np.asarray(tri, dtype=np.float64)
    a, b, c = t[:, 0], t[:, 1], t[:, 2]
    area = 0.5 * np.linalg.norm(np.cross(b - a, c - a), axis=1).sum()
    vol = np.einsum("ij,ij->i", a, np.cross(b, c)).sum() / 6.0
    return area, vol


def open_edges(tri, decimals=None):
    """Number of undirected edges not shared by exactly two triangles (0 for a closed surface).  Vertices are matched
    bit-exactly (the kernel derives every shared vertex from the same end point)."""
    t = np.asarray(tri, dtype=np.float32).reshape(-1, 3)
    keys = t.view([("x", "f4"), ("y", "f4"), ("z", "f4")]).reshape(-1)
    _, ids = np.unique(keys, return_inverse=True)
    f = ids.reshape(-1, 3)
    f = f[(f[:, 0] != f[:, 1]) & (f[:, 1] != f[:, 2]) & (f[:, 0] != f[:, 2])]          # drop degenerate (zero-area) triangles
    e = np.concatenate([f[:, [0, 1]], f[:, [1, 2]], f[:, [2, 0]]], 0)
    e.sort(axis=1)
    _, cnt = np.unique(e, axis=0, return_counts=True)
    return int((cnt != 2).sum())
