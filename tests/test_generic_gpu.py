"""GPU parity of the drop-in Decoder (generic fp32 CUDA path, through the C ABI) against the golden vectors of the
REAL reference and against the CPU oracle.  Tolerances: fp32 path, 1e-5 abs on SDF, 1e-4 relative on gradients."""
import ctypes as C
import os

import numpy as np
import pytest
import torch

from deepsdf_b200 import _lib as L
from oracle import decoder_oracle as O
from tests.util import GOLDEN, build_exact_module as build_module, load_cases

pytestmark = pytest.mark.gpu
META, BLOB = load_cases()
DEV = "cuda:0"


def _inputs(name):
    m = META[name]
    pre = name + "::"
    table = torch.from_numpy(BLOB[pre + "table"].copy())
    rows = torch.from_numpy(np.repeat(BLOB[pre + "scenes"], m["sps"])).long()
    xyz = torch.from_numpy(BLOB[pre + "xyz"])
    gt = torch.from_numpy(BLOB[pre + "gt"])
    return m, pre, table, rows, xyz, gt


def test_reference_export_outputs():
    """The reference's own trained + exported 4x64 network: identical outputs (fp32 path)."""
    z = np.load(os.path.join(GOLDEN, "small_net_export.npz"))
    params = {k[len("param::"):]: torch.from_numpy(z[k]) for k in z.files if k.startswith("param::")}
    mod = build_module(META["small_4x64_L2_tanh"]["spec"], params, DEV).eval()
    with torch.no_grad():
        y = mod(torch.from_numpy(z["input"]).to(DEV)).cpu().numpy()
    assert np.abs(y - z["output"]).max() <= 1e-5


@pytest.mark.parametrize("name", sorted(META))
def test_eval_forward_and_input_grad(name):
    m, pre, table, rows, xyz, gt = _inputs(name)
    ospec = O.OracleSpec(**m["spec"])
    params = O.make_params(ospec, m["seed"])
    mod = build_module(m["spec"], params, DEV).eval()
    inp = torch.cat([table[rows], xyz], 1).to(DEV).requires_grad_(True)
    y = mod(inp)
    assert y.shape == (inp.shape[0], 1)
    assert np.abs(y.detach().cpu().numpy() - BLOB[pre + "eval_out"]).max() <= 1e-5
    (G,) = torch.autograd.grad(y.sum(), inp)
    ref_G = BLOB[pre + "input_grad"]
    assert np.abs(G.cpu().numpy() - ref_G).max() <= 1e-4 * max(1.0, np.abs(ref_G).max())


@pytest.mark.parametrize("name", sorted(META))
def test_train_step_grads_match_reference(name):
    """train_deep_sdf.py:483-533 re-typed around the drop-in module (decoder.eval(): no RNG), against the golden
    loss / latent grad / parameter grads produced by the real reference."""
    m, pre, table, rows, xyz, gt = _inputs(name)
    ospec = O.OracleSpec(**m["spec"])
    params = O.make_params(ospec, m["seed"])
    mod = build_module(m["spec"], params, DEV).eval()
    lat = torch.nn.Embedding(table.shape[0], table.shape[1], max_norm=m["code_bound"])
    with torch.no_grad():
        lat.weight.copy_(table)
    clampd, lam, epoch = m["clamp"], m["lam"], m["epoch"]
    n_total = xyz.shape[0]
    sdf_gt = torch.clamp(gt, -clampd, clampd)
    batch_vecs = lat(rows)
    inp = torch.cat([batch_vecs, xyz], 1)
    pred = mod(inp.to(DEV))
    pred_c = torch.clamp(pred, -clampd, clampd)
    loss = torch.nn.L1Loss(reduction="sum")(pred_c, sdf_gt.to(DEV)) / n_total
    reg = (lam * min(1, epoch / 100) * torch.sum(torch.norm(batch_vecs, dim=1))) / n_total
    loss = loss + reg.to(DEV)
    loss.backward()
    assert abs(float(loss) - float(BLOB[pre + "loss"][0])) <= 2e-6
    ref_lg = BLOB[pre + "lat_grad"]
    assert np.abs(lat.weight.grad.numpy() - ref_lg).max() <= 1e-6 + 2e-4 * np.abs(ref_lg).max()
    for k, p in mod.named_parameters():
        g = p.grad.detach().cpu().numpy()
        ref_norm = float(BLOB[pre + "gnorm::" + k][0])
        assert abs(np.linalg.norm(g.astype(np.float64)) - ref_norm) <= 2e-4 * ref_norm + 1e-8, k
        if m["big"]:
            idx = BLOB[pre + "gidx::" + k]
            rg = BLOB[pre + "gsample::" + k]
            assert np.abs(g.reshape(-1)[idx] - rg).max() <= 1e-6 + 2e-3 * np.abs(rg).max(), k
        else:
            rg = BLOB[pre + "grad::" + k]
            assert np.abs(g - rg).max() <= 1e-6 + 2e-4 * np.abs(rg).max(), k


@pytest.mark.parametrize("name", ["small_4x64_L2_tanh", "xyz_in_all_4x40_L4", "layernorm_4x48_L3", "big_8x512_L16"])
def test_training_mode_with_shared_dropout_masks(name):
    """Dropout parity needs a shared mask: the CUDA path exports its Philox keep-masks (dsdf_dropout_mask) and the
    oracle consumes them."""
    m, pre, table, rows, xyz, gt = _inputs(name)
    kw = m["spec"]
    ospec = O.OracleSpec(**kw)
    ospec.dropout_scale = O.kernel_dropout_scale(float(ospec.dropout_prob))   # shared masks: the kernels' survivor scale
    params = O.make_params(ospec, m["seed"])
    mod = build_module(kw, params, DEV).train()
    inp_cpu = torch.cat([table[rows], xyz], 1)
    N = inp_cpu.shape[0]
    torch.manual_seed(77)
    seed = int(torch.randint(0, 2 ** 62, (1,)).item())
    torch.manual_seed(77)
    inp = inp_cpu.to(DEV).requires_grad_(True)
    y = mod(inp)
    w = torch.linspace(-1, 1, N, device=DEV).reshape(-1, 1)
    (y * w).sum().backward()
    lib = L.lib()
    masks = {}
    for l in range(ospec.num_linear - 1):
        if ospec.dropout is not None and l in ospec.dropout and ospec.dropout_prob > 0:
            buf = torch.empty(N, ospec.out_dims[l], dtype=torch.uint8, device=DEV)
            L.check(lib.dsdf_dropout_mask(seed, l, 0, N, ospec.out_dims[l], float(ospec.dropout_prob), buf.data_ptr(), None))
            masks[l] = buf.cpu().float()
            frac = 1.0 - masks[l].mean().item()
            assert abs(frac - lib.dsdf_dropout_effective_p(float(ospec.dropout_prob))) < 0.03
    lat_mask = None
    if ospec.latent_dropout:
        buf = torch.empty(N, ospec.latent_size, dtype=torch.uint8, device=DEV)
        L.check(lib.dsdf_dropout_mask(seed, 0xFFFF, 0, N, ospec.latent_size, 0.2, buf.data_ptr(), None))
        lat_mask = buf.cpu().float()
    torch.cuda.synchronize()
    P = {k: v.clone().requires_grad_(True) for k, v in params.items()}
    x = inp_cpu.clone().requires_grad_(True)
    yo = O.decoder_forward(ospec, P, x, training=True, masks=masks, latent_mask=lat_mask)
    (yo * w.cpu()).sum().backward()
    assert (y.detach().cpu() - yo.detach()).abs().max().item() <= 1e-5
    assert (inp.grad.cpu() - x.grad).abs().max().item() <= 1e-4 * max(1.0, x.grad.abs().max().item())
    for k, p in mod.named_parameters():
        ref = P[k].grad
        assert (p.grad.cpu() - ref).abs().max().item() <= 1e-6 + 2e-4 * ref.abs().max().item(), k


def test_jvp_double_backward_matches_G():
    """deep_sdf/mesh.py:405-422 computes torch.autograd.functional.jvp through the decoder (double-backward trick)."""
    name = "small_4x32_L1"
    m, pre, table, rows, xyz, gt = _inputs(name)
    ospec = O.OracleSpec(**m["spec"])
    params = O.make_params(ospec, m["seed"])
    mod = build_module(m["spec"], params, DEV).eval()
    inp_cpu = torch.cat([table[rows], xyz], 1)
    G = O.input_gradient(ospec, params, inp_cpu)
    lat = inp_cpu[:, 0].to(DEV)
    rest = inp_cpu[:, 1:].to(DEV)
    v = torch.linspace(0.5, 1.5, inp_cpu.shape[0], device=DEV)

    def f(latcol):
        return mod(torch.cat([latcol.reshape(-1, 1), rest], 1)).squeeze(1)

    out, jv = torch.autograd.functional.jvp(f, lat, v=v)
    ref = G[:, 0] * v.cpu()
    assert (jv.cpu() - ref).abs().max().item() <= 1e-5 * max(1.0, ref.abs().max().item())


def test_dataparallel_wrapper_and_cpu_input():
    """train_deep_sdf.py:353 always wraps in DataParallel and hands it a CPU tensor (lat_vecs lives on the CPU)."""
    name = "small_4x64_L2_tanh"
    m, pre, table, rows, xyz, gt = _inputs(name)
    ospec = O.OracleSpec(**m["spec"])
    params = O.make_params(ospec, m["seed"])
    mod = build_module(m["spec"], params).cuda()
    dp = torch.nn.DataParallel(mod, device_ids=[0]).eval()
    inp = torch.cat([table[rows], xyz], 1)
    y = dp(inp)
    assert y.is_cuda and np.abs(y.detach().cpu().numpy() - BLOB[pre + "eval_out"]).max() <= 1e-5
    sd = dp.state_dict()
    assert all(k.startswith("module.lin") for k in sd)


def test_chunked_recompute_path(monkeypatch):
    import deepsdf_b200.networks.deep_sdf_decoder as D
    monkeypatch.setattr(D, "_WS_SINGLE_PASS_LIMIT", 1)
    monkeypatch.setattr(D, "_CHUNK_ROWS", 40)
    name = "small_4x32_L1"
    m, pre, table, rows, xyz, gt = _inputs(name)
    ospec = O.OracleSpec(**m["spec"])
    params = O.make_params(ospec, m["seed"])
    mod = build_module(m["spec"], params, DEV).eval()
    inp = torch.cat([table[rows], xyz], 1).to(DEV).requires_grad_(True)
    y = mod(inp)
    y.sum().backward()
    assert np.abs(y.detach().cpu().numpy() - BLOB[pre + "eval_out"]).max() <= 1e-5
    P = {k: v.clone().requires_grad_(True) for k, v in params.items()}
    yo = O.decoder_forward(ospec, P, inp.detach().cpu())
    yo.sum().backward()
    for k, p in mod.named_parameters():
        assert (p.grad.cpu() - P[k].grad).abs().max().item() <= 1e-6 + 2e-4 * P[k].grad.abs().max().item(), k


def test_dropin_decoder_in_reference_loop_reproduces_main_function_fixture():
    """BASELINE.json config #1 (scaled down): the drop-in Decoder used exactly as train_deep_sdf.py:346-545 uses the
    reference class (DataParallel wrapper, nn.Embedding(max_norm) on the CPU, L1Loss(sum)/N, code regulariser,
    torch.optim.Adam over both groups) replays the losses and final parameters that the REAL main_function produced
    (tests/golden/train_epochs.npz, written by oracle/make_golden_epoch.py)."""
    from tests.test_oracle_golden import load_train_epochs
    z, specs, kw, init, final = load_train_epochs()
    decoder = torch.nn.DataParallel(build_module(kw, init, DEV))                       # :346-353
    lat_vecs = torch.nn.Embedding(6, specs["CodeLength"], max_norm=specs["CodeBound"])  # :385
    with torch.no_grad():
        lat_vecs.weight.copy_(torch.from_numpy(z["lat_init"]))
    loss_l1 = torch.nn.L1Loss(reduction="sum")                                         # :398
    sched = specs["LearningRateSchedule"]
    opt = torch.optim.Adam([{"params": decoder.parameters(), "lr": sched[0]["Initial"]},
                            {"params": lat_vecs.parameters(), "lr": sched[1]["Initial"]}])
    sps, ipe, clampd = specs["SamplesPerScene"], int(z["its_per_epoch"]), specs["ClampingDistance"]
    decoder.train()
    for it in range(len(z["loss"])):
        epoch = it // ipe + 1
        sdf_data = torch.from_numpy(z["batches"][it])
        n = sdf_data.shape[0]
        xyz, sdf_gt = sdf_data[:, 0:3], torch.clamp(sdf_data[:, 3].unsqueeze(1), -clampd, clampd)
        indices = torch.from_numpy(z["indices"][it]).unsqueeze(-1).repeat(1, sps).view(-1)
        opt.zero_grad()
        batch_vecs = lat_vecs(indices)
        pred = torch.clamp(decoder(torch.cat([batch_vecs, xyz], dim=1)), -clampd, clampd)     # CPU tensor in, like :511-514
        loss = loss_l1(pred, sdf_gt.to(DEV)) / n
        reg = specs["CodeRegularizationLambda"] * min(1, epoch / 100) * torch.sum(torch.norm(batch_vecs, dim=1)) / n
        loss = loss + reg.to(DEV)
        loss.backward()
        opt.step()
        ref = float(z["loss"][it])
        assert abs(loss.item() - ref) <= 1e-5 * abs(ref), (it, loss.item(), ref)
    sd = decoder.module.state_dict()
    for k, v in final.items():
        assert torch.allclose(sd[k].cpu(), v, rtol=1e-4, atol=5e-6), k
    assert torch.allclose(lat_vecs.weight.detach(), torch.from_numpy(z["lat_final"]), rtol=1e-4, atol=5e-6)


def test_create_graph_backward_gives_parameter_gradients_and_guards_second_order():
    """loss.backward(create_graph=True) must still deliver first-order parameter gradients (it returned None), and a
    second-order request through the kernel's input gradient must raise instead of silently contributing zero."""
    m, pre, table, rows, xyz, gt = _inputs("small_4x64_L2_tanh")
    ospec = O.OracleSpec(**m["spec"])
    params = O.make_params(ospec, m["seed"])
    mod = build_module(m["spec"], params, DEV).eval()
    inp = torch.cat([table[rows], xyz], 1).to(DEV).requires_grad_(True)
    w = torch.linspace(-1, 1, inp.shape[0], device=DEV).reshape(-1, 1)
    # reference gradients (plain backward)
    (mod(inp) * w).sum().backward()
    ref_p = {k: p.grad.clone() for k, p in mod.named_parameters()}
    ref_x = inp.grad.clone()
    mod.zero_grad()
    inp.grad = None
    # create_graph=True: same first-order gradients
    (mod(inp) * w).sum().backward(create_graph=True)
    for k, p in mod.named_parameters():
        assert p.grad is not None and torch.allclose(p.grad.detach(), ref_p[k], rtol=1e-5, atol=1e-7), k
    assert torch.allclose(inp.grad.detach(), ref_x, rtol=1e-5, atol=1e-7)
    mod.zero_grad()
    inp.grad = None
    # an eikonal-style term: d/dparams of |dSDF/dx| needs second-order kernels -> loud failure
    y = mod(inp)
    (gx,) = torch.autograd.grad(y.sum(), inp, create_graph=True)
    pen = (gx[:, -3:].norm(dim=1) - 1).pow(2).mean()
    with pytest.raises(RuntimeError, match="second-order"):
        pen.backward()


def test_forward_without_dropout_leaves_the_rng_alone():
    """a training-mode forward of a spec WITHOUT dropout must not draw from torch's global generator (the reference's
    DataLoader shuffle shares it)."""
    m = META["geom2d_3x24_L2"]
    mod = build_module(m["spec"], None, DEV).train()
    x = torch.randn(64, 4, device=DEV)
    torch.manual_seed(123)
    a = torch.rand(3)
    torch.manual_seed(123)
    mod(x)
    b = torch.rand(3)
    assert torch.equal(a, b)


def test_jit_trace_records_the_registered_op(tmp_path):
    """create_libtorch_executable.py:19-20 on a CUDA example input: torch.jit.trace of the drop-in module records the
    torch.library op deepsdf_b200::decoder_eval; the traced / scripted / saved + reloaded module reproduces the module."""
    m, pre, table, rows, xyz, gt = _inputs("small_4x64_L2_tanh")
    ospec = O.OracleSpec(**m["spec"])
    params = O.make_params(ospec, m["seed"])
    mod = build_module(m["spec"], params, DEV).eval()
    inp = torch.cat([table[rows], xyz], 1).to(DEV)
    with torch.no_grad():
        want = mod(inp)
        traced = torch.jit.trace(mod, inp[:1])
    assert "deepsdf_b200::decoder_eval" in str(traced.graph)
    sm = torch.jit.script(traced)
    fn = str(tmp_path / "cpp_model.pt")
    sm.save(fn)
    again = torch.jit.load(fn)
    with torch.no_grad():
        assert torch.equal(traced(inp), want) and torch.equal(again(inp), want)
