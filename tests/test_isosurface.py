"""Iso-surface extraction (the tail of create_mesh, deep_sdf/mesh.py:88-155).  skimage's marching cubes is absent here and
the kernel is marching tetrahedra, so: the CPU restatement is checked against analytic geometry (CPU test), the kernel is
checked bit-exactly against the restatement and, at create_mesh's N = 256, against the analytic sphere (GPU tests)."""
import os

import numpy as np
import pytest
import torch

from oracle import isosurface_oracle as IO


def _sphere(n, r=0.6, c=(0.05, -0.1, 0.02)):
    g = np.linspace(-1, 1, n, dtype=np.float32)
    x, y, z = np.meshgrid(g, g, g, indexing="ij")
    return (np.sqrt((x - c[0]) ** 2 + (y - c[1]) ** 2 + (z - c[2]) ** 2) - r).astype(np.float32)


def _check_sphere(tri, n, r=0.6, c=(0.05, -0.1, 0.02), rel=0.02):
    h = 2.0 / (n - 1)
    d = np.linalg.norm(tri.reshape(-1, 3).astype(np.float64) - np.asarray(c), axis=1)
    assert np.abs(d - r).max() <= h * h / r + 1e-6                       # linear interpolation error of a curved level set
    area, vol = IO.mesh_area_volume(tri)
    assert abs(area - 4 * np.pi * r * r) <= rel * 4 * np.pi * r * r
    assert abs(vol - 4 / 3 * np.pi * r ** 3) <= rel * 4 / 3 * np.pi * r ** 3     # positive: normals point outwards


def test_oracle_marching_tetrahedra_on_a_sphere():
    n = 20
    tri = IO.marching_tetrahedra(_sphere(n), (-1, -1, -1), [2.0 / (n - 1)] * 3)
    assert tri.shape[0] > 1000
    _check_sphere(tri, n, rel=0.03)
    assert IO.open_edges(tri) == 0                                       # closed surface
    assert IO.marching_tetrahedra(np.ones((5, 5, 5), np.float32)).shape == (0, 3, 3)


@pytest.mark.gpu
def test_kernel_matches_oracle_bit_exactly():
    from deepsdf_b200.mesh import extract_isosurface
    rng = np.random.RandomState(3)
    cases = [(_sphere(14), (-1, -1, -1), 2.0 / 13, 0.0)]
    g = np.linspace(-1, 1, 11, dtype=np.float32)
    x, y, z = np.meshgrid(g[:9], g, g[:7], indexing="ij")                # non-cubic volume (a z-slab), non-zero level
    cases.append(((np.sin(3 * x) * np.cos(2 * y) + 0.5 * z + 0.1 * rng.normal(size=x.shape)).astype(np.float32), (0.5, -2.0, 1.0),
                  [0.1, 0.2, 0.3], 0.05))
    for vol, origin, voxel, level in cases:
        want = IO.marching_tetrahedra(vol, origin, voxel if isinstance(voxel, list) else [voxel] * 3, level)
        got = extract_isosurface(torch.from_numpy(vol).cuda(), origin, voxel, level).cpu().numpy()
        assert got.shape == want.shape and want.shape[0] > 100
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32))


@pytest.mark.gpu
def test_sphere_at_create_mesh_resolution_and_ply(tmp_path):
    from deepsdf_b200.mesh import convert_sdf_samples_to_ply, extract_isosurface
    n = 256
    vol = torch.from_numpy(_sphere(n)).cuda()
    tri = extract_isosurface(vol, (-1, -1, -1), 2.0 / (n - 1)).cpu().numpy()
    _check_sphere(tri, n, rel=0.002)
    assert IO.open_edges(tri) == 0
    assert extract_isosurface(torch.ones(8, 8, 8, device="cuda")).shape == (0, 3, 3)          # nothing to extract
    fn = str(tmp_path / "sphere.ply")
    n_tri = convert_sdf_samples_to_ply(vol, [-1, -1, -1], 2.0 / (n - 1), fn, offset=np.array([0.1, 0.0, 0.0]), scale=2.0)
    assert n_tri == tri.shape[0]
    raw = open(fn, "rb").read()
    head, body = raw.split(b"end_header\n", 1)
    assert b"element vertex %d" % (3 * n_tri) in head and b"element face %d" % n_tri in head
    assert len(body) == 3 * n_tri * 12 + n_tri * 13
    v = np.frombuffer(body[:3 * n_tri * 12], dtype="<f4").reshape(-1, 3)
    np.testing.assert_allclose(v, tri.reshape(-1, 3) / 2.0 - np.array([0.1, 0, 0], np.float32), rtol=0, atol=1e-6)   # mesh.py:123-127
