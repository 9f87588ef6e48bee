"""Pins oracle/decoder_oracle.py against vectors produced by the REAL reference
(oracle/make_golden.py): the reference's TorchScript export cpp_model.pt and its Decoder
class + train-step math + torch.optim.Adam on nine NetworkSpecs."""
import json
import os

import numpy as np
import pytest
import torch

from oracle import decoder_oracle as O

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")
torch.set_num_threads(1)


def _cases():
    with open(os.path.join(GOLDEN, "decoder_cases.json")) as f:
        return json.load(f)


CASES = _cases()
BLOB = np.load(os.path.join(GOLDEN, "decoder_cases.npz"))


def test_small_net_export_matches_torchscript_outputs():
    z = np.load(os.path.join(GOLDEN, "small_net_export.npz"))
    params = {k[len("param::"):]: torch.from_numpy(z[k]) for k in z.files if k.startswith("param::")}
    spec = O.OracleSpec(latent_size=2, dims=[64] * 4, geom_dimension=3, dropout=list(range(8)), dropout_prob=0.2,
                        norm_layers=list(range(8)), latent_in=[1], weight_norm=True, xyz_in_all=False, use_tanh=True)
    assert sorted(params) == sorted(spec.param_names())
    for k, shp in spec.param_shapes().items():
        assert tuple(params[k].shape) == shp
    y = O.decoder_forward(spec, params, torch.from_numpy(z["input"]), training=False)
    assert np.abs(y.numpy() - z["output"]).max() <= 1e-6


@pytest.mark.parametrize("name", sorted(CASES))
def test_oracle_matches_reference_decoder(name):
    m = CASES[name]
    spec = O.OracleSpec(**m["spec"])
    params = O.make_params(spec, m["seed"])
    pre = name + "::"
    table = torch.from_numpy(BLOB[pre + "table"].copy())
    scenes = BLOB[pre + "scenes"]
    rows = torch.from_numpy(np.repeat(scenes, m["sps"])).long()
    xyz = torch.from_numpy(BLOB[pre + "xyz"])
    gt = torch.from_numpy(BLOB[pre + "gt"])
    inp = torch.cat([table[rows], xyz], 1)
    y = O.decoder_forward(spec, params, inp, training=False)
    assert np.abs(y.numpy() - BLOB[pre + "eval_out"]).max() <= 5e-6
    G = O.input_gradient(spec, params, inp)
    ref_G = BLOB[pre + "input_grad"]
    assert np.abs(G.numpy() - ref_G).max() <= 1e-5 * max(1.0, np.abs(ref_G).max())

    loss, grads, lat_grad, pred = O.train_loss_and_grads(
        spec, params, table, rows, xyz, gt, m["clamp"], m["lam"], m["epoch"], training=False,
        code_bound=m["code_bound"])
    assert abs(loss - float(BLOB[pre + "loss"][0])) <= 1e-6
    assert np.abs(table.numpy() - BLOB[pre + "table_after_renorm"]).max() <= 1e-6
    assert np.abs(pred.numpy() - BLOB[pre + "pred_train"]).max() <= 5e-6
    ref_lg = BLOB[pre + "lat_grad"]
    assert np.abs(lat_grad.numpy() - ref_lg).max() <= 1e-6 + 1e-4 * np.abs(ref_lg).max()
    # Adam (both groups) after one step
    lr_dec, lr_lat = m["lr"]
    for k in spec.param_names():
        g = grads[k].numpy()
        ref_norm = float(BLOB[pre + "gnorm::" + k][0])
        assert abs(np.linalg.norm(g.astype(np.float64)) - ref_norm) <= 1e-4 * ref_norm + 1e-9
        p = params[k].clone()
        O.adam_step_(p, grads[k], torch.zeros_like(p), torch.zeros_like(p), 1, lr_dec)
        if m["big"]:
            idx = BLOB[pre + "gidx::" + k]
            rg = BLOB[pre + "gsample::" + k]
            assert np.abs(g.reshape(-1)[idx] - rg).max() <= 1e-6 + 1e-3 * np.abs(rg).max()
            assert np.abs(p.numpy().reshape(-1)[idx] - BLOB[pre + "postsample::" + k]).max() <= 2e-6
        else:
            rg = BLOB[pre + "grad::" + k]
            assert np.abs(g - rg).max() <= 1e-6 + 1e-4 * np.abs(rg).max()
            # Adam's first step is lr*sign(g) wherever |g| >> eps; compare away from g ~ 0
            sel = np.abs(rg) > 1e-6
            assert np.abs(p.numpy() - BLOB[pre + "post::" + k])[sel].max(initial=0.0) <= 2e-6
    zl = table.clone()
    O.adam_step_(zl, lat_grad, torch.zeros_like(zl), torch.zeros_like(zl), 1, lr_lat)
    sel = np.abs(ref_lg) > 1e-6
    assert np.abs(zl.numpy() - BLOB[pre + "post_lat"])[sel].max(initial=0.0) <= 2e-6


def test_grid_points_layout():
    # deep_sdf/mesh.py:37-53 : index = (i*N + j)*N + k, coord = -1 + idx*2/(N-1)
    N = 5
    p = O.grid_points(N)
    assert p.shape == (N ** 3, 3)
    i, j, k = 3, 1, 4
    row = (i * N + j) * N + k
    assert np.allclose(p[row].numpy(), [-1 + i * 0.5, -1 + j * 0.5, -1 + k * 0.5])


def test_analytic_sdfs():
    x = np.array([[0.0, 0.0, 0.0], [1.0, 1.0, 1.0], [0.5, 0.0, 0.0]])
    s = O.corner_spheres_sdf(x, 0.5, limit=0.9)
    assert s.shape == (3, 1) and abs(s[0, 0] - (-0.9)) < 1e-12 and s[1, 0] > 0
    c = O.cross_ms_sdf(x, 0.3)
    assert abs(c[0, 0] - (-0.3)) < 1e-12


# ---------------------------------------------------------------------------------------------------------------
# BASELINE.json config #1 (scaled down): two epochs of the REAL train_deep_sdf.main_function, recorded by
# oracle/make_golden_epoch.py -> tests/golden/train_epochs.npz.  The oracle's training loop must reproduce it.
def load_train_epochs():
    import json
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", "train_epochs.npz"))
    specs = json.loads(str(z["specs_json"]))
    kw = dict(latent_size=specs["CodeLength"], **specs["NetworkSpecs"])
    init = {k[len("init::"):]: torch.from_numpy(z[k]) for k in z.files if k.startswith("init::")}
    final = {k[len("final::"):]: torch.from_numpy(z[k]) for k in z.files if k.startswith("final::")}
    return z, specs, kw, init, final


def test_oracle_training_loop_reproduces_reference_main_function():
    z, specs, kw, init, final = load_train_epochs()
    spec = O.OracleSpec(**kw)
    assert list(init) == spec.param_names()                     # the reference's own parameter names and order
    P = {k: v.clone() for k, v in init.items()}
    m1 = {k: torch.zeros_like(v) for k, v in P.items()}
    m2 = {k: torch.zeros_like(v) for k, v in P.items()}
    Z = torch.from_numpy(z["lat_init"]).clone()
    zm, zv = torch.zeros_like(Z), torch.zeros_like(Z)
    sps, ipe = specs["SamplesPerScene"], int(z["its_per_epoch"])
    sched = specs["LearningRateSchedule"]
    for it in range(len(z["loss"])):
        epoch = it // ipe + 1
        batch = torch.from_numpy(z["batches"][it])
        row_scene = torch.from_numpy(z["indices"][it]).repeat_interleave(sps)
        loss, grads, zgrad, _ = O.train_loss_and_grads(spec, P, Z, row_scene, batch[:, :3], batch[:, 3], specs["ClampingDistance"],
                                                       specs["CodeRegularizationLambda"], epoch, training=True,
                                                       code_bound=specs["CodeBound"])
        assert abs(loss - float(z["loss"][it])) <= 2e-6 * abs(float(z["loss"][it])), (it, loss, float(z["loss"][it]))
        lr_d = O.step_lr(sched[0]["Initial"], sched[0]["Interval"], sched[0]["Factor"], epoch)
        lr_z = O.step_lr(sched[1]["Initial"], sched[1]["Interval"], sched[1]["Factor"], epoch)
        for k in P:
            O.adam_step_(P[k], grads[k], m1[k], m2[k], it + 1, lr_d)
        O.adam_step_(Z, zgrad, zm, zv, it + 1, lr_z)
    for k in P:
        assert torch.allclose(P[k], final[k], rtol=1e-5, atol=2e-6), k
    assert torch.allclose(Z, torch.from_numpy(z["lat_final"]), rtol=1e-5, atol=2e-6)
