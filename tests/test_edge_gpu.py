"""Edge cases and size-independent properties of the drop-in ``Decoder`` and the fused path on the GPU:
empty and ragged inputs (the reference's ``nn.Linear`` chain accepts any N, including 0), single rows, chunk invariance
(rows are independent: evaluating a batch in pieces gives the same bits), run-to-run determinism of the training
gradients at BASELINE.json's config #2 size, and z-slab consistency of the grid evaluation."""
import numpy as np
import pytest
import torch

from oracle import decoder_oracle as O
from tests.util import build_exact_module as build_module, load_cases, sdf_scale_params

pytestmark = pytest.mark.gpu
META, BLOB = load_cases()
DEV = "cuda:0"


def _module(name, out_scale=0.025):
    kw = dict(META[name]["spec"])
    ospec = O.OracleSpec(**kw)
    params = sdf_scale_params(ospec, META[name]["seed"], out_scale=out_scale)
    return kw, ospec, params, build_module(kw, params, DEV).eval()


@pytest.mark.parametrize("name", ["big_8x512_L16", "small_4x64_L2_tanh"])
def test_empty_and_ragged_inputs(name):
    kw, ospec, params, mod = _module(name)
    width = ospec.latent_size + 3
    y0 = mod(torch.zeros(0, width, device=DEV))
    assert y0.shape == (0, 1) and y0.dtype == torch.float32                     # what an nn.Linear chain returns for N = 0
    rng = np.random.RandomState(5)
    for n in (1, 2, 127, 128, 129, 1000):
        x = torch.from_numpy(np.concatenate([rng.normal(0, 0.1, (n, ospec.latent_size)), rng.uniform(-1, 1, (n, 3))], 1).astype(np.float32))
        y = mod(x.to(DEV)).detach().cpu()
        want = O.decoder_forward(ospec, params, x, training=False)
        assert y.shape == (n, 1)
        tol = 2e-3 if name.startswith("big") else 2e-5                           # bf16 tensor-core chain / fp32 generic path
        assert float((y - want).abs().max()) <= tol, (n, float((y - want).abs().max()))


def test_rows_are_independent_chunk_invariance():
    """Same bits whether 1000 rows go through in one call or in ragged pieces (no cross-row state, no tile-position
    dependence); also after a row permutation."""
    kw, ospec, params, mod = _module("big_8x512_L16")
    rng = np.random.RandomState(6)
    n = 1000
    x = torch.from_numpy(np.concatenate([rng.normal(0, 0.1, (n, 16)), rng.uniform(-1, 1, (n, 3))], 1).astype(np.float32)).to(DEV)
    whole = mod(x)
    pieces = torch.cat([mod(x[a:b]) for a, b in ((0, 1), (1, 130), (130, 517), (517, 1000))], 0)
    assert torch.equal(whole, pieces)
    perm = torch.from_numpy(rng.permutation(n)).to(DEV)
    assert torch.equal(mod(x[perm]), whole[perm])


def test_training_step_is_deterministic_at_config2_size():
    """BASELINE.json config #2 (64 scenes x 16384 rows, 8x512, L = 256): two runs of the fused training step from the
    same state give bit-identical losses, decoder gradients and latent gradients (fixed-order reductions, no atomics),
    and the loss equals the sum of |clamp(pred) - clamp(gt)| / N recomputed from the kernel's own SDF output."""
    import bench
    from deepsdf_b200.trainer import SceneTrainer
    out = []
    for _ in range(2):
        tr = SceneTrainer(bench.SPECS, bench.SCENES_TOTAL, torch.device(DEV), seed=0)
        gen = torch.Generator(device=DEV).manual_seed(2)
        ids = torch.arange(64) * 3
        batch = bench.make_batch(gen, torch.device(DEV), ids, bench.SCENES_TOTAL)
        loss = tr.step(batch, ids.to(DEV), epoch=1).clone()
        sdf = tr.engine._bufs["sdf"].clone()
        out.append((loss, tr.flat_g.clone(), tr.lat_grad.clone(), sdf, batch))
        del tr
        torch.cuda.empty_cache()
    (l0, g0, z0, s0, b0), (l1, g1, z1, s1, _) = out
    assert torch.equal(l0, l1) and torch.equal(g0, g1) and torch.equal(z0, z1) and torch.equal(s0, s1)
    assert torch.isfinite(g0).all() and float(g0.abs().max()) > 0
    n = b0.shape[0]
    data_term = ((s0.clamp(-0.1, 0.1) - b0[:, 3].clamp(-0.1, 0.1)).abs().double().sum() / n)
    assert abs(float(l0[1]) - float(data_term)) <= 2e-6 * max(1.0, float(data_term)), (float(l0[1]), float(data_term))


def test_grid_slabs_tile_the_volume():
    """create_mesh's grid in z-slabs (one per rank at N GPUs) equals the one-piece evaluation, values and gradients."""
    from deepsdf_b200.mesh import create_mesh_sdf
    from deepsdf_b200.sharding import slab_for_rank
    kw, ospec, params, mod = _module("big_8x512_L16")
    lat = torch.randn(16) * 0.1
    N = 40
    whole, dxyz, dlat = create_mesh_sdf(mod, lat, N=N, with_grad=True, max_batch=4096)
    parts = [create_mesh_sdf(mod, lat, N=N, with_grad=True, slab=slab_for_rank(N, r, 3), max_batch=8192) for r in range(3)]
    assert torch.equal(torch.cat([p[0] for p in parts], 0), whole)
    assert torch.equal(torch.cat([p[1] for p in parts], 0), dxyz)
    assert torch.equal(torch.cat([p[2] for p in parts], 0), dlat)
    assert whole.shape == (N, N, N) and dxyz.shape == (N, N, N, 3) and dlat.shape == (N, N, N, 16)


def test_grid_512_slabs_equal_whole_volume():
    """BASELINE.json config #4 size: the 512^3 grid evaluated as 8 z-slabs (what 8 GPUs do) is bit-identical to the
    one-piece evaluation, and the grid points are the reference's (mesh.py:37-53) at the volume's corners."""
    from deepsdf_b200.mesh import create_mesh_sdf
    from deepsdf_b200.sharding import slab_for_rank
    kw, ospec, params, mod = _module("big_8x512_L16")
    lat = torch.randn(16, generator=torch.Generator().manual_seed(4)) * 0.1
    N = 512
    whole = create_mesh_sdf(mod, lat, N=N)
    assert whole.shape == (N, N, N) and bool(torch.isfinite(whole).all())
    for r in range(8):
        i0, i1 = slab_for_rank(N, r, 8)
        part = create_mesh_sdf(mod, lat, N=N, slab=(i0, i1), max_batch=1 << 20)
        assert torch.equal(part, whole[i0:i1]), r
    corners = torch.tensor([[-1.0, -1.0, -1.0], [1.0, 1.0, 1.0], [-1.0, 1.0, -1.0]])
    q = torch.cat([lat.expand(3, -1), corners], 1)
    want = O.decoder_forward(ospec, params, q, training=False).squeeze(1)
    got = torch.stack([whole[0, 0, 0], whole[-1, -1, -1], whole[0, -1, 0]]).cpu()
    assert float((got - want).abs().max()) <= 2e-3
