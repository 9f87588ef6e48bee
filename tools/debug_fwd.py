#!/usr/bin/env python3
"""Layer-by-layer comparison of the fused forward's stashed activations with the CPU bf16 emulation."""
import ctypes as C
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from deepsdf_b200 import _lib as L  # noqa: E402
from deepsdf_b200.spec import NetSpec  # noqa: E402
from oracle import decoder_oracle as O  # noqa: E402
from tests.util import emulate_fused_forward, sdf_scale_params, bf16r  # noqa: E402
from tools.bringup import to_params_struct  # noqa: E402

dev = torch.device("cuda:0")
lib = L.lib()
st = lambda: L.stream_ptr(dev)


def run(name, n_rows=1000, p_drop=0.0):
    meta = json.load(open(os.path.join(ROOT, "tests", "golden", "decoder_cases.json")))[name]
    kw = dict(meta["spec"])
    kw["dropout_prob"] = p_drop
    ospec = O.OracleSpec(**kw)
    spec = NetSpec(**kw)
    params_cpu = sdf_scale_params(ospec, meta["seed"])
    params = {k: v.to(dev).contiguous() for k, v in params_cpu.items()}
    desc = spec.desc()
    nbytes = lib.dsdf_pack_bytes(C.byref(desc))
    pack = torch.zeros(nbytes + 1024, dtype=torch.uint8, device=dev)
    off = (-pack.data_ptr()) % 1024
    pack = pack[off:off + nbytes]
    P = to_params_struct(spec, params)
    L.check(lib.dsdf_pack_weights(C.byref(desc), C.byref(P), pack.data_ptr(), st()))
    rng = np.random.RandomState(3)
    n_tiles = (n_rows + 127) // 128
    n_pad = n_tiles * 128
    nh = spec.num_linear - 1
    table = torch.from_numpy((rng.normal(0, 1, size=(5, spec.latent_size)) / np.sqrt(spec.latent_size)).astype(np.float32))
    scene_ids = torch.tensor([3, 0, 4], dtype=torch.int64)
    tile_scene = torch.from_numpy(rng.randint(0, 3, size=n_tiles).astype(np.int32))
    xyz = torch.from_numpy(rng.uniform(-1, 1, size=(n_rows, 3)).astype(np.float32))
    gt = torch.from_numpy(rng.normal(0, 0.08, size=(n_rows,)).astype(np.float32))
    cbias = torch.zeros(3, 2, 512, device=dev)
    table_d = table.to(dev)
    L.check(lib.dsdf_scene_prepare(C.byref(desc), pack.data_ptr(), table_d.data_ptr(), scene_ids.to(dev).data_ptr(), 3, 0.0,
                                   cbias.data_ptr(), st()))
    out = torch.full((n_rows,), float("nan"), device=dev)
    acts = torch.full((nh, n_pad, 512), float("nan"), dtype=torch.bfloat16, device=dev)
    bitmask = torch.zeros(nh, n_pad, 16, dtype=torch.int32, device=dev)
    dy = torch.zeros(n_pad, device=dev)
    tile_loss = torch.zeros(n_tiles, device=dev)
    io = L.TrainIO()
    io.b_last = params[f"lin{spec.num_linear - 1}.bias"].data_ptr()
    io.training = 1
    io.seed = 1234
    io.row0 = 0
    gt_d = gt.to(dev)
    io.gt = gt_d.data_ptr()
    io.clamp = 0.1
    io.inv_n = 1.0 / n_rows
    io.dy_out = dy.data_ptr()
    io.tile_loss = tile_loss.data_ptr()
    io.bitmask = bitmask.data_ptr()
    io.acts = acts.data_ptr()
    io.n_rows_pad = n_pad
    xyz_d, ts_d = xyz.to(dev), tile_scene.to(dev)
    L.check(lib.dsdf_fused_forward(C.byref(desc), pack.data_ptr(), cbias.data_ptr(), ts_d.data_ptr(), xyz_d.data_ptr(), 3,
                                   n_rows, out.data_ptr(), C.byref(io), st()))
    torch.cuda.synchronize()
    row_scene = scene_ids[tile_scene.long()].repeat_interleave(128)[:n_rows]
    masks = None
    if p_drop > 0:
        masks = {}
        for l in range(nh):
            buf = torch.empty(n_rows, ospec.out_dims[l], dtype=torch.uint8, device=dev)
            L.check(lib.dsdf_dropout_mask(1234, l, 0, n_rows, ospec.out_dims[l], p_drop, buf.data_ptr(), None))
            masks[l] = buf.cpu().float()
    emu, y, eacts = emulate_fused_forward(ospec, params_cpu, table, row_scene, xyz, masks=masks)
    # cbias slot 1
    ll = [l for l in range(1, nh) if l in ospec.latent_in]
    if ll:
        l4 = ll[0]
        W4 = O.effective_weight(ospec, params_cpu, l4)
        xw = ospec.in_dims[l4] - spec.latent_size - 3
        c4 = table[scene_ids] @ W4[:, xw:xw + spec.latent_size].t() + params_cpu[f"lin{l4}.bias"]
        print("cbias slot1 err", (cbias[:, 1, :].cpu() - c4).abs().max().item())
    for l in range(nh):
        w = ospec.out_dims[l]
        got = acts[l, :n_rows, :w].float().cpu()
        want = bf16r(eacts[l])
        d = (got - want).abs()
        rel = d.max().item() / max(want.abs().max().item(), 1e-9)
        nbad = int((d > 1e-2 * want.abs().max()).sum())
        # bitmask check
        bits = bitmask[l, :n_rows].cpu().numpy().view(np.uint32)
        bm = ((bits[:, :, None] >> np.arange(32)[None, None, :]) & 1).reshape(n_rows, 512)[:, :w]
        bm_ok = (bm == (want.numpy() > 0)).mean()
        print(f"layer {l}: act max_abs_diff={d.max().item():.4g} (rel {rel:.3g}) mean={d.mean().item():.3g} n_bad={nbad} "
              f"want_max={want.abs().max().item():.3g} nan={int(torch.isnan(got).sum())} bitmask_agree={bm_ok:.6f}")
    e = (out.cpu() - emu.squeeze(1)).abs()
    print(f"out vs emulation: max {e.max().item():.3g} mean {e.mean().item():.3g}; rows>1e-4: {int((e > 1e-4).sum())}/{n_rows}")
    ref = O.decoder_forward(ospec, params_cpu, torch.cat([table[row_scene], xyz], 1), training=p_drop > 0, masks=masks).squeeze(1)
    e2 = (out.cpu() - ref).abs()
    print(f"out vs fp32 oracle: max {e2.max().item():.3g} mean {e2.mean().item():.3g}")
    e3 = (emu.squeeze(1) - ref).abs()
    print(f"emulation vs fp32 oracle: max {e3.max().item():.3g} mean {e3.mean().item():.3g}")
    # loss / dy
    t = emu.squeeze(1)
    pc, gc = t.clamp(-0.1, 0.1), gt.clamp(-0.1, 0.1)
    loss_ref = (pc - gc).abs().sum().item()
    print(f"tile_loss sum {tile_loss.sum().item():.6f} vs emu {loss_ref:.6f}")
    dt = (1 - t * t)
    dy_ref = torch.sign(pc - gc) * ((t >= -0.1) & (t <= 0.1)).float() * dt / n_rows
    print("dy err", (dy[:n_rows].cpu() - dy_ref).abs().max().item(), "dy max", dy_ref.abs().max().item())


if __name__ == "__main__":
    for nm in sys.argv[1:] or ["big_8x512_L256"]:
        print("=====", nm, "no dropout")
        run(nm, p_drop=0.0)
        print("=====", nm, "dropout 0.2")
        run(nm, p_drop=0.2)
