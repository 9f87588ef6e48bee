"""The split-fp16 high-precision eval forward (csrc/chain_exact.cu, ``FusedEngine.decode_scenes(exact=True)``) against the
fp32 arithmetic of the reference Decoder (deep_sdf/networks/deep_sdf_decoder.py:76-111): the north star's 1e-5 bound on the
tensor cores.  Checked on the three 8x512 network classes of the golden cases at NATURAL output scale (|sdf| up to ~1, no
shrunken last layer), against a float64 evaluation of the oracle (so that the comparison is not limited by the fp32 CPU
reference's own rounding), with ragged row counts, several scenes and a renormalised code; the bf16 chain runs beside it
on the same inputs to show what the mode buys."""
import numpy as np
import pytest
import torch

from oracle import decoder_oracle as O
from tests.util import build_module, load_cases

pytestmark = pytest.mark.gpu
META, BLOB = load_cases()
DEV = "cuda:0"
BIG = ["big_8x512_L256", "big_8x512_L2", "big_8x512_L16"]
TOL = 1e-5


def _f64(params):
    return {k: v.double() for k, v in params.items()}


@pytest.mark.parametrize("name", BIG)
@pytest.mark.parametrize("tail", [0, 37, 100])
def test_exact_forward_within_1e5_of_fp32_arithmetic(name, tail):
    from deepsdf_b200.engine import FusedEngine
    kw = dict(META[name]["spec"])
    ospec = O.OracleSpec(**kw)
    params = O.make_params(ospec, META[name]["seed"])                 # default-initialised network: outputs span (-1, 1)
    mod = build_module(kw, params, DEV).eval()
    eng = FusedEngine(mod)
    rng = np.random.RandomState(5)
    Lz = ospec.latent_size
    table = torch.from_numpy((rng.normal(0, 1, (6, Lz)) / np.sqrt(Lz)).astype(np.float32))
    scene_ids = torch.tensor([4, 1, 2, 0, 5], dtype=torch.int64)
    tiles_per_scene = 3
    tile_scene = torch.arange(len(scene_ids), dtype=torch.int32).repeat_interleave(tiles_per_scene)
    n = tile_scene.numel() * 128 - tail
    xyz = torch.from_numpy(rng.uniform(-1, 1, (n, 3)).astype(np.float32))
    row_scene = scene_ids[tile_scene.long()].repeat_interleave(128)[:n]
    cb = eng.scene_bias(table.to(DEV), scene_ids.to(DEV))
    cbx = eng.scene_bias_exact(table.to(DEV), scene_ids.to(DEV))
    y_exact = eng.decode_scenes(cbx, tile_scene.to(DEV), xyz.to(DEV), exact=True).cpu()
    y_bf16 = eng.decode_scenes(cb, tile_scene.to(DEV), xyz.to(DEV)).cpu()
    inp = torch.cat([table[row_scene], xyz], 1)
    ref64 = O.decoder_forward(ospec, _f64(params), inp.double()).squeeze(1)
    ref32 = O.decoder_forward(ospec, params, inp).squeeze(1)
    e_exact = (y_exact.double() - ref64).abs()
    e_bf16 = (y_bf16.double() - ref64).abs()
    e_ref32 = (ref32.double() - ref64).abs()
    print(f"{name} n={n}: |sdf| max {float(ref64.abs().max()):.3f}; exact mode max {float(e_exact.max()):.2e} mean {float(e_exact.mean()):.2e}; "
          f"bf16 chain max {float(e_bf16.max()):.2e}; fp32 CPU reference vs float64 max {float(e_ref32.max()):.2e}")
    assert not torch.isnan(y_exact).any()
    assert float(e_exact.max()) <= TOL
    assert float(e_exact.max()) < 0.05 * float(e_bf16.max())          # it is a different precision class, not a lucky draw


def test_exact_pack_follows_parameter_updates_and_grid_path():
    """create_mesh_sdf(exact=True): the 64^3 grid of one code; a parameter update between calls re-packs the (hi, lo) weights."""
    from deepsdf_b200.mesh import create_mesh_sdf
    name = "big_8x512_L16"
    kw = dict(META[name]["spec"])
    ospec = O.OracleSpec(**kw)
    params = O.make_params(ospec, 3)
    mod = build_module(kw, params, DEV).eval()
    lat = torch.from_numpy(np.random.RandomState(2).normal(0, 0.25, (ospec.latent_size,)).astype(np.float32))
    NG = 32
    for rep in range(2):
        vol = create_mesh_sdf(mod, lat, N=NG, max_batch=20000, exact=True)
        idx = torch.arange(NG ** 3)
        vs = 2.0 / (NG - 1)
        grid = torch.stack([(idx // (NG * NG)) * vs - 1, ((idx // NG) % NG) * vs - 1, (idx % NG) * vs - 1], 1).float()
        q = torch.cat([lat.expand(NG ** 3, -1), grid], 1)
        cur = {k: v.detach().cpu().clone() for k, v in mod.state_dict().items()}
        ref = O.decoder_forward(ospec, _f64(cur), q.double()).squeeze(1)
        err = (vol.reshape(-1).cpu().double() - ref).abs().max().item()
        print(f"grid {NG}^3 exact mode, pass {rep}: max err {err:.2e}")
        assert err <= TOL
        with torch.no_grad():
            for p in mod.parameters():
                p.mul_(1.01)


def test_decode_sdf_honours_the_modules_precision_switch():
    """``decoder.fused_exact = True`` routes decode_sdf(decoder, latent, queries) (deep_sdf/utils.py:54-65) and create_mesh's grid
    to the high-precision mode without touching the reference's call signatures."""
    from deepsdf_b200.utils import decode_sdf
    name = "big_8x512_L256"
    kw = dict(META[name]["spec"])
    ospec = O.OracleSpec(**kw)
    params = O.make_params(ospec, 7)
    mod = build_module(kw, params, DEV).eval()
    rng = np.random.RandomState(4)
    lat = torch.from_numpy((rng.normal(0, 1, (ospec.latent_size,)) / 16).astype(np.float32))
    pts = torch.from_numpy(rng.uniform(-1, 1, (1000, 3)).astype(np.float32))
    ref = O.decoder_forward(ospec, _f64(params), torch.cat([lat.expand(1000, -1), pts], 1).double())
    y_bf16 = decode_sdf(mod, lat.to(DEV), pts.to(DEV)).cpu()
    mod.fused_exact = True
    y_exact = decode_sdf(mod, lat.to(DEV), pts.to(DEV)).cpu()
    mod.fused_exact = False
    e_b, e_x = (y_bf16.double() - ref).abs().max().item(), (y_exact.double() - ref).abs().max().item()
    print(f"decode_sdf: bf16 chain max err {e_b:.2e}, fused_exact max err {e_x:.2e}")
    assert y_exact.shape == (1000, 1) and e_x <= TOL and e_x < 0.05 * e_b
