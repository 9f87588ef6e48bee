"""Small launches of every hand-written kernel family, sized for `make sanitize` (compute-sanitizer memcheck / racecheck /
synccheck run them 10-100x slower): one fused training step (forward, dgrad chain, wgrad, finish, Adam, pack), a forward over
more tiles than CTA pairs... scaled down to a few tiles per pair, the input-gradient consumers, the small-network kernel, the
sampler and the iso-surface.  They also run in the normal GPU suite (seconds)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
NET = dict(dims=[512] * 8, dropout=list(range(8)), dropout_prob=0.2, norm_layers=list(range(8)), latent_in=[4],
           weight_norm=True, xyz_in_all=False, use_tanh=False, latent_dropout=False, geom_dimension=3)
SPECS = {"NetworkSpecs": NET, "CodeLength": 16,
         "LearningRateSchedule": [{"Type": "Step", "Initial": 5e-4, "Interval": 500, "Factor": 0.5},
                                  {"Type": "Step", "Initial": 1e-3, "Interval": 500, "Factor": 0.5}],
         "SamplesPerScene": 384, "ScenesPerBatch": 3, "ClampingDistance": 0.1, "CodeRegularization": True,
         "CodeRegularizationLambda": 1e-4, "CodeBound": 1.0}


def test_training_steps_small():
    from deepsdf_b200.trainer import SceneTrainer
    tr = SceneTrainer(SPECS, 5, DEV, seed=2)
    g = torch.Generator().manual_seed(0)
    batch = torch.cat([torch.rand(3 * 384, 3, generator=g) * 2 - 1, torch.randn(3 * 384, 1, generator=g) * 0.05], 1).to(DEV)   # 9 tiles: odd
    ids = torch.tensor([0, 2, 4], device=DEV)
    losses = [float(tr.step(batch, ids, epoch=3)[0]) for _ in range(3)]
    tr.batch_split = 2
    losses.append(float(tr.step(batch, ids, epoch=3)[0]))
    torch.cuda.synchronize()
    assert all(np.isfinite(losses))


def test_eval_forward_and_input_gradient_small():
    from deepsdf_b200.networks.deep_sdf_decoder import Decoder
    from deepsdf_b200.utils import fused_engine_for
    torch.manual_seed(1)
    for L in (16, 2):
        dec = Decoder(L, **NET).to(DEV).eval()
        eng = fused_engine_for(dec)
        n = 5 * 128
        lat = torch.randn(1, L, device=DEV) * 0.2
        cb = eng.scene_bias(lat, torch.zeros(1, dtype=torch.int64, device=DEV))
        xyz = torch.rand(n, 3, device=DEV) * 2 - 1
        ts = torch.zeros(n // 128, dtype=torch.int32, device=DEV)
        sdf = eng.decode_scenes(cb, ts, xyz)
        sdf2, G = eng.sdf_and_input_grad(cb, ts, xyz)
        if L == 2:
            q = torch.cat([lat.expand(n, -1), xyz], 1).contiguous()
            s3, G3 = eng.sdf_and_input_grad_rows(q)
            assert torch.isfinite(G3).all()
        torch.cuda.synchronize()
        assert torch.isfinite(sdf).all() and torch.isfinite(G).all() and float((sdf - sdf2).abs().max()) <= 1e-3


def test_small_network_sampler_isosurface_small():
    from deepsdf_b200.data import DeviceSDFSamples
    from deepsdf_b200.mesh import extract_isosurface
    from deepsdf_b200.networks.deep_sdf_decoder import Decoder
    from deepsdf_b200.small import small_engine_for
    torch.manual_seed(2)
    dec = Decoder(2, dims=[64] * 4, geom_dimension=3, dropout=[0, 1, 2, 3], dropout_prob=0.2, norm_layers=[0, 1, 2, 3, 4], latent_in=[1],
                  weight_norm=True, xyz_in_all=False, use_tanh=True, latent_dropout=False).to(DEV).eval()
    sdf, G = small_engine_for(dec).sdf_and_input_grad(torch.randn(300, 5, device=DEV) * 0.5)
    rng = np.random.RandomState(0)
    arrays = []
    for s in range(3):
        rows = np.concatenate([rng.uniform(-1, 1, (900, 3)), rng.normal(0, 0.1, (900, 1))], 1).astype(np.float32)
        arrays.append((rows[rows[:, 3] > 0], rows[rows[:, 3] <= 0]))
    ds = DeviceSDFSamples(None, None, 256, 3, DEV, arrays=arrays)
    b = ds.sample_batch(torch.tensor([2, 0], device=DEV), 5)
    gl = torch.linspace(-1, 1, 24, device=DEV)
    vol = (torch.sqrt(gl[:, None, None] ** 2 + gl[None, :, None] ** 2 + gl[None, None, :] ** 2) - 0.6).contiguous()
    tri = extract_isosurface(vol, (-1, -1, -1), 2.0 / 23)
    torch.cuda.synchronize()
    assert torch.isfinite(sdf).all() and torch.isfinite(G).all() and b.shape == (512, 4) and tri.shape[0] > 100
