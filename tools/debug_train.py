#!/usr/bin/env python3
"""GPU diagnostics of the fused training step: per-parameter gradient errors vs the CPU oracle, split by column block."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from deepsdf_b200 import _lib as L  # noqa: E402
from oracle import decoder_oracle as O  # noqa: E402
from tests.util import build_module, load_cases, sdf_scale_params  # noqa: E402
from tests.test_fused_gpu import _batch  # noqa: E402

META, _ = load_cases()
DEV = "cuda:0"


def rel(a, b):
    return float((a - b).norm() / (b.norm() + 1e-20))


def run(name, drop):
    from deepsdf_b200.engine import FusedEngine
    kw = dict(META[name]["spec"]); kw["dropout_prob"] = drop
    ospec = O.OracleSpec(**kw)
    params = sdf_scale_params(ospec, META[name]["seed"], out_scale=0.03)
    mod = build_module(kw, params, DEV)
    eng = FusedEngine(mod)
    table, scene_ids, tile_scene, xyz, gt, row_scene = _batch(ospec)
    N = xyz.shape[0]
    clampd, lam_eff, bound, seed = 0.1, 1e-4 * 0.37, 1.0, 4242
    masks = {}
    if drop > 0:
        for l in range(ospec.num_linear - 1):
            buf = torch.empty(N, ospec.out_dims[l], dtype=torch.uint8, device=DEV)
            L.check(eng.lib.dsdf_dropout_mask(seed, l, 0, N, ospec.out_dims[l], drop, buf.data_ptr(), None))
            masks[l] = buf.cpu().float()
    table_o = table.clone()
    O.embedding_renorm_(table_o, row_scene, bound)
    xyz, gt, bad = O.well_conditioned_batch(ospec, params, table_o, row_scene, masks if drop > 0 else None, clampd, seed=5)
    tab_d = table.clone().to(DEV)
    grads, PG = eng.alloc_grads()
    lat_grad = torch.zeros_like(tab_d)
    loss = eng.loss_and_grads(tab_d, scene_ids.to(DEV), tile_scene.to(DEV), xyz.to(DEV), gt.to(DEV), clamp=clampd,
                              reg_lambda_eff=lam_eff, n_total=N, code_bound=bound, seed=seed, row0=0, PG=PG, lat_grad=lat_grad)
    torch.cuda.synchronize()
    ref_loss, ref_grads, ref_lat, pred = O.train_loss_and_grads(ospec, params, table.clone(), row_scene, xyz, gt, clampd, lam_eff * 100, 1,
                                                                masks=masks if drop > 0 else None, code_bound=bound)
    print(f"== {name} drop={drop} unsafe={bad} loss {float(loss[0]):.6f} vs {ref_loss:.6f}")
    sdf = eng._bufs["sdf"].cpu()
    print(f"   sdf vs oracle pred: max {float((sdf - pred.squeeze(1)).abs().max()):.2e}")
    dy = eng._bufs["dy"].view(torch.float32)[:N].cpu()
    # oracle dy: dL/d(pre-tanh)
    p = pred.squeeze(1)
    inside = (p.abs() <= clampd).float()
    dy_ref = torch.sign(p.clamp(-clampd, clampd) - gt.clamp(-clampd, clampd)) * inside * (1 - p * p) / N
    print(f"   dy rel err {rel(dy, dy_ref):.3e}  flips {(torch.sign(dy) != torch.sign(dy_ref)).sum().item()}")
    names = mod.flat_param_names()
    Lz, D = ospec.latent_size, 3
    for nm, g in zip(names, grads):
        r = ref_grads[nm]
        g = g.cpu()
        extra = ""
        if nm.endswith("original1") and (nm.startswith("lin0.") or nm.startswith("lin4.")):
            xw = r.shape[1] - (Lz + D)
            parts = []
            if xw > 0:
                parts.append(("x", slice(0, xw)))
            parts += [("lat", slice(xw, xw + Lz)), ("xyz", slice(xw + Lz, xw + Lz + D))]
            extra = " | " + " ".join(f"{t}:{rel(g[:, sl], r[:, sl]):.2e}" for t, sl in parts)
        print(f"   {nm:48s} {rel(g, r):.3e} |ref| {float(r.norm()):.2e}{extra}")
    print(f"   latent grad {rel(lat_grad.cpu(), ref_lat):.3e}")


if __name__ == "__main__":
    run("big_8x512_L16", 0.0)
    run("big_8x512_L256", 0.2)
