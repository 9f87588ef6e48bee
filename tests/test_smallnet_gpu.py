"""The fused fp32 small-network kernel (csrc/smallnet.cu) against the golden vectors of the REAL reference Decoder
(tests/golden/decoder_cases.*: eval outputs and input gradients of the 4x64 / 4x32 / 6x128 networks), against the
reference's own trained + exported network (experiments/corner_spheres_only_small_network/cpp_model.pt ->
tests/golden/small_net_export.npz), against the CPU oracle on ragged sizes, and through the consumers that route to it
(decode_sdf, create_mesh_sdf, evaluate_network_sdf - deep_sdf/utils.py:54-65, mesh.py:37-73, mesh.py:457-528).
Tolerances: fp32 path, 1e-5 abs on SDF, 1e-4 relative on gradients (the north star's high-precision bound)."""
import os

import numpy as np
import pytest
import torch

from deepsdf_b200 import _lib as L
from oracle import decoder_oracle as O
from tests.util import GOLDEN, build_module, load_cases

pytestmark = pytest.mark.gpu
META, BLOB = load_cases()
DEV = "cuda:0"
SMALL = ["small_4x64_L2_tanh", "small_4x32_L1", "mid_6x128_L1"]


def _case(name):
    m = META[name]
    pre = name + "::"
    table = torch.from_numpy(BLOB[pre + "table"].copy())
    rows = torch.from_numpy(np.repeat(BLOB[pre + "scenes"], m["sps"])).long()
    xyz = torch.from_numpy(BLOB[pre + "xyz"])
    ospec = O.OracleSpec(**m["spec"])
    params = O.make_params(ospec, m["seed"])
    mod = build_module(m["spec"], params, DEV).eval()
    return m, pre, ospec, params, mod, torch.cat([table[rows], xyz], 1)


def _grad_close(G, refG, what):
    """G is the gradient of a piecewise-linear function: a row whose pre-activation sits within fp32 summation noise of a ReLU
    kink takes the other branch under a different (equally valid) summation order, and its gradient jumps.  Such rows are
    rare (a handful per million gates); every other row must agree to 1e-4 relative."""
    tol = 1e-4 * max(1.0, refG.abs().max().item())
    bad = ((G - refG).abs().amax(dim=1) > tol)
    assert bad.float().mean().item() <= 2e-3 and int(bad.sum()) <= 3 + G.shape[0] // 500, (what, int(bad.sum()), G.shape[0])
    return int(bad.sum())


def test_support_matrix():
    from deepsdf_b200.small import small_engine_for
    for name in sorted(META):
        kw = META[name]["spec"]
        mod = build_module(kw, None, DEV)
        eng = small_engine_for(mod)
        want = max(kw["dims"]) <= 128 and kw["weight_norm"] in (True, False) and not kw["xyz_in_all"] and not kw["latent_dropout"] \
            and not (not kw["weight_norm"] and len(kw["norm_layers"]) > 0)
        assert (eng is not None) == want, name


@pytest.mark.parametrize("name", SMALL)
def test_golden_eval_output_and_input_gradient(name):
    from deepsdf_b200.small import small_engine_for
    m, pre, ospec, params, mod, inp = _case(name)
    eng = small_engine_for(mod)
    sdf, G = eng.sdf_and_input_grad(inp.to(DEV))
    torch.cuda.synchronize()
    e_out = np.abs(sdf.cpu().numpy() - BLOB[pre + "eval_out"].reshape(-1)).max()
    ref_G = BLOB[pre + "input_grad"]
    e_G = np.abs(G.cpu().numpy() - ref_G).max()
    print(f"{name}: vs the real reference Decoder: sdf max err {e_out:.2e}, G max err {e_G:.2e} (|G| max {np.abs(ref_G).max():.2f})")
    assert e_out <= 1e-5
    assert e_G <= 1e-4 * max(1.0, np.abs(ref_G).max())
    # forward-only launch gives the same SDF bit for bit
    sdf2, none = eng.sdf_and_input_grad(inp.to(DEV), with_grad=False)
    assert none is None and torch.equal(sdf2, sdf)


def test_reference_trained_export():
    """The reference's own trained 4x64 network (cpp_model.pt): same outputs on its recorded inputs."""
    from deepsdf_b200.small import small_engine_for
    z = np.load(os.path.join(GOLDEN, "small_net_export.npz"))
    params = {k[len("param::"):]: torch.from_numpy(z[k]) for k in z.files if k.startswith("param::")}
    mod = build_module(META["small_4x64_L2_tanh"]["spec"], params, DEV).eval()
    sdf, _ = small_engine_for(mod).sdf_and_input_grad(torch.from_numpy(z["input"]).to(DEV), with_grad=False)
    assert np.abs(sdf.cpu().numpy() - z["output"].reshape(-1)).max() <= 1e-5


@pytest.mark.parametrize("name", SMALL)
def test_ragged_sizes_broadcast_latent_and_repack(name):
    """Row counts that are not tile multiples (1, 127, 129, 1000, 5 tiles + 3), a broadcast code, and a parameter update
    between calls (the cached pack must follow torch's version counters)."""
    from deepsdf_b200.small import small_engine_for
    m, pre, ospec, params, mod, _ = _case(name)
    eng = small_engine_for(mod)
    g = torch.Generator().manual_seed(4)
    Lz = ospec.latent_size
    for n in (1, 127, 129, 1000, 5 * 128 + 3):
        q = torch.cat([torch.randn(n, Lz, generator=g) * 0.4, torch.rand(n, 3, generator=g) * 2 - 1], 1)
        sdf, G = eng.sdf_and_input_grad(q.to(DEV))
        ref = O.decoder_forward(ospec, params, q).squeeze(1)
        refG = O.input_gradient(ospec, params, q)
        assert (sdf.cpu() - ref).abs().max().item() <= 1e-5, n
        _grad_close(G.cpu(), refG, n)
    lat = torch.randn(Lz, generator=g) * 0.3
    pts = torch.rand(777, 3, generator=g) * 2 - 1
    sdf, G = eng.sdf_and_input_grad(pts.to(DEV), latent=lat.to(DEV))
    q = torch.cat([lat.expand(777, -1), pts], 1)
    assert (sdf.cpu() - O.decoder_forward(ospec, params, q).squeeze(1)).abs().max().item() <= 1e-5
    _grad_close(G.cpu(), O.input_gradient(ospec, params, q), "broadcast latent")
    # in-place parameter update -> next call re-packs
    with torch.no_grad():
        mod.lin0.bias.add_(0.05)
    params2 = {k: v.detach().cpu().clone() for k, v in mod.state_dict().items()}
    sdf3, _ = small_engine_for(mod).sdf_and_input_grad(q.to(DEV), with_grad=False)
    ref3 = O.decoder_forward(ospec, params2, q).squeeze(1)
    assert (sdf3.cpu() - ref3).abs().max().item() <= 1e-5
    assert (ref3 - O.decoder_forward(ospec, params, q).squeeze(1)).abs().max().item() > 1e-4      # the update mattered


def test_consumers_route_small_networks():
    """decode_sdf / create_mesh_sdf / evaluate_network_sdf on the networks the reference's optimisation runs use
    (optimization_runs/*/config.json:8): 4x32 (double_lattice_3D_small_network) and 6x128 (round_cross_big_network)."""
    from deepsdf_b200.mesh import create_mesh_sdf, evaluate_network_sdf, location_lookup
    from deepsdf_b200.utils import decode_sdf
    for name in ("small_4x32_L1", "mid_6x128_L1", "small_4x64_L2_tanh"):
        m, pre, ospec, params, mod, inp = _case(name)
        Lz = ospec.latent_size
        rng = np.random.RandomState(3)
        # decode_sdf, both forms (deep_sdf/utils.py:54-65)
        lat = torch.from_numpy(rng.normal(0, 0.3, (Lz,)).astype(np.float32))
        pts = torch.from_numpy(rng.uniform(-1, 1, (999, 3)).astype(np.float32))
        y = decode_sdf(mod, lat.to(DEV), pts.to(DEV))
        ref = O.decoder_forward(ospec, params, torch.cat([lat.expand(999, -1), pts], 1))
        assert y.shape == (999, 1) and (y.cpu() - ref).abs().max().item() <= 1e-5
        y2 = decode_sdf(mod, None, inp[:500].to(DEV))
        assert (y2.cpu() - O.decoder_forward(ospec, params, inp[:500])).abs().max().item() <= 1e-5
        # create_mesh grid with derivatives (mesh.py:37-73 + 405-454)
        NG = 20
        vol, dxyz, dlat = create_mesh_sdf(mod, lat, N=NG, max_batch=3000, with_grad=True)
        idx = torch.arange(NG ** 3)
        vs = 2.0 / (NG - 1)
        grid = torch.stack([(idx // (NG * NG)) * vs - 1, ((idx // NG) % NG) * vs - 1, (idx % NG) * vs - 1], 1).float()
        q = torch.cat([lat.expand(NG ** 3, -1), grid], 1)
        assert (vol.reshape(-1).cpu() - O.decoder_forward(ospec, params, q).squeeze(1)).abs().max().item() <= 1e-5
        refG = O.input_gradient(ospec, params, q)
        _grad_close(torch.cat([dlat.reshape(-1, Lz).cpu(), dxyz.reshape(-1, 3).cpu()], 1), refG, "grid")
        # the microstructure evaluation with per-vertex codes, tiling and caps (mesh.py:457-514 restated around the oracle)
        n, tiling = 3001, [2, 3, 1]
        samples_orig = torch.from_numpy((rng.uniform(-0.5, 0.5, (n, 3)) * 2.1).astype(np.float32))
        lv = torch.from_numpy(rng.normal(0, 0.4, (n, Lz)).astype(np.float32))
        caps = {"x0": {"cap": 1, "measure": 0.1}, "z1": {"cap": -1, "measure": 0.05}}
        sdf, dl = evaluate_network_sdf(mod, lv.to(DEV), samples_orig.to(DEV), tiling, caps, with_grad=True)
        samples = samples_orig.clone()
        for d, t in enumerate(tiling):                                                    # mesh.py:470-477
            p = 2.0 / t
            samples[:, d] = (2.0 / p) * torch.abs(torch.remainder(samples[:, d] - (t % 2), 2 * p) - p) - 1
        qv = torch.cat([lv, samples], 1)
        s_ref = O.decoder_forward(ospec, params, qv).squeeze(1)
        G_ref = O.input_gradient(ospec, params, qv)[:, :Lz]
        capped = torch.zeros(n, dtype=torch.bool)
        for loc, cd in caps.items():                                                       # mesh.py:499-512
            dim, mult = location_lookup[loc]
            border = (samples_orig[:, dim] - mult * (1 - cd["measure"])) * -mult
            new = torch.maximum(s_ref, -border) if cd["cap"] == -1 else torch.minimum(s_ref, border)
            capped |= new != s_ref
            s_ref = new
        outside = (samples_orig.abs() - 1).amax(dim=1)                                      # unit-cube cap, mesh.py:513-514
        new = torch.maximum(s_ref, outside)
        capped |= new != s_ref
        s_ref = new
        assert (sdf.cpu() - s_ref).abs().max().item() <= 1e-5, name
        free = ~capped
        _grad_close(dl.cpu()[free], G_ref[free], "evaluate_network")
        assert (dl.cpu()[capped] == 0).all()


@pytest.mark.parametrize("name", ["small_4x32_L1", "mid_6x128_L1"])
def test_latent_optimizer_on_small_networks_tracks_oracle_reconstruct(name):
    """LatentOptimizer (upstream reconstruct.py, batched over shapes) on the small-network kernel: fp32 arithmetic, so the loss
    curve follows the oracle's torch loop closely (the clamped-L1 sign gradient still amplifies last-bit differences)."""
    from deepsdf_b200.reconstruct import LatentOptimizer
    m, pre, ospec, params, mod, _ = _case(name)
    Lz = ospec.latent_size
    rng = np.random.RandomState(21)
    S, M, ns, iters = 3, 1500, 200, 30            # ragged: 200 rows per shape, shapes not tile aligned
    z_true = torch.from_numpy(rng.normal(0, 0.5, (S, Lz)).astype(np.float32))
    xyz = torch.from_numpy(rng.uniform(-1, 1, (S, M, 3)).astype(np.float32))
    with torch.no_grad():
        gt = torch.stack([0.2 * O.decode_sdf(ospec, params, z_true[s], xyz[s]).squeeze(1) for s in range(S)])
    data = torch.cat([xyz, gt.unsqueeze(-1)], -1)
    idx = torch.from_numpy(rng.randint(0, M, (iters, S, ns)))
    init = torch.from_numpy(rng.normal(0, 0.01, (S, Lz)).astype(np.float32))
    opt = LatentOptimizer(mod, S, num_samples=ns, init=init)
    assert opt.small is not None
    hist = opt.reconstruct(data.to(DEV), num_iterations=iters, sample_idx=idx, record_every=1)
    got_loss = torch.stack(hist).cpu()
    for s in range(S):
        z = init[s:s + 1].clone().requires_grad_(True)
        o = torch.optim.Adam([z], lr=5e-3)
        losses = []
        for e in range(iters):
            for pg in o.param_groups:
                pg["lr"] = 5e-3 * (0.1 ** (e // int(iters / 2)))
            q, g = xyz[s][idx[e, s]], torch.clamp(gt[s][idx[e, s]].reshape(-1, 1), -0.1, 0.1)
            o.zero_grad()
            pred = torch.clamp(O.decoder_forward(ospec, params, torch.cat([z.expand(ns, -1), q], 1)), -0.1, 0.1)
            loss = (pred - g).abs().mean() + 1e-4 * torch.mean(z.pow(2))
            loss.backward()
            o.step()
            losses.append(float(loss.detach()))
        ref_loss = torch.tensor(losses)
        rel = ((got_loss[:, s] - ref_loss).abs() / ref_loss).max().item()
        print(f"{name} shape {s}: loss {ref_loss[0]:.5f} -> {ref_loss[-1]:.5f}; max rel loss diff {rel:.2e}")
        assert rel <= 5e-3 and abs(float(got_loss[0, s]) - float(ref_loss[0])) <= 1e-5 * float(ref_loss[0]) + 1e-7


@pytest.mark.parametrize("name", SMALL)
def test_boundary_call_routes_eval_mode_small_networks(name):
    """decoder(input[N, L+3]) itself - what the reference's unchanged scripts call - in eval mode: one small-network launch for
    the SDF (+ G when the input requires grad), the jvp double-backward trick of deep_sdf/mesh.py:405-422, and parameter
    gradients (their own graph branch, fp32 generic path) equal to the ones with the routing switched off."""
    m, pre, ospec, params, mod, inp = _case(name)
    lib = L.lib()
    x = inp.to(DEV)
    with torch.no_grad():
        n0 = int(lib.dsdf_launch_count())
        y = mod(x)
        n_launch = int(lib.dsdf_launch_count()) - n0
    assert y.shape == (x.shape[0], 1) and n_launch <= 2, n_launch
    assert np.abs(y.squeeze(1).cpu().numpy() - BLOB[pre + "eval_out"].reshape(-1)).max() <= 1e-5
    xg = x.clone().requires_grad_(True)
    n0 = int(lib.dsdf_launch_count())
    yg = mod(xg)
    (G,) = torch.autograd.grad(yg.sum(), xg)
    assert int(lib.dsdf_launch_count()) - n0 <= 2                      # SDF and G from the same launch, no parameter-gradient pass
    ref_G = BLOB[pre + "input_grad"]
    assert np.abs(G.cpu().numpy() - ref_G).max() <= 1e-4 * max(1.0, np.abs(ref_G).max())
    # forward-mode product through the double-backward trick
    v = torch.randn_like(x)
    _, jv = torch.autograd.functional.jvp(lambda t: mod(t), x, v)
    want = (torch.from_numpy(ref_G).to(DEV) * v).sum(1, keepdim=True)
    assert (jv - want).abs().max().item() <= 1e-4 * max(1.0, float(want.abs().max()))
    # parameter gradients: same values as the generic path
    mod.zero_grad()
    mod(x).sum().backward()
    g_small = [p.grad.clone() for p in mod.parameters()]
    mod.fused_eval = False
    mod.zero_grad()
    y32 = mod(x)
    y32.sum().backward()
    mod.fused_eval = True
    assert all(torch.equal(a, p.grad) for a, p in zip(g_small, mod.parameters()))
    assert (y32.detach() - y).abs().max().item() <= 1e-5
