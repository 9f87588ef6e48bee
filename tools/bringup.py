#!/usr/bin/env python3
"""GPU bring-up diagnostics (run on the B200 box): tcgen05 descriptor self-tests with convention sweeps and a first
fused-forward parity check.  Prints one line per experiment; never raises on numerical mismatch."""
import ctypes as C
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from deepsdf_b200 import _lib as L  # noqa: E402
from deepsdf_b200.spec import NetSpec  # noqa: E402
from oracle import decoder_oracle as O  # noqa: E402
from tests.util import emulate_fused_forward, sdf_scale_params  # noqa: E402

dev = torch.device("cuda:0")
lib = L.lib()
L.check(lib.dsdf_check_device())
st = lambda: L.stream_ptr(dev)


def selftest(mode, K, N, a_lbo, a_sbo, b_lbo, b_sbo, tag):
    g = torch.Generator(device="cpu").manual_seed(K * 1000 + N + mode)
    a_mn, b_mn = mode & 1, mode & 2
    A = torch.randn(128, K, generator=g).to(torch.bfloat16)      # logical [M][K]
    B = torch.randn(N, K, generator=g).to(torch.bfloat16)        # logical [N][K]
    ref = A.float() @ B.float().t()
    A_dev = (A.t().contiguous() if a_mn else A.contiguous()).to(dev)
    B_dev = (B.t().contiguous() if b_mn else B.contiguous()).to(dev)
    Cout = torch.full((128, N), float("nan"), device=dev)
    store = torch.zeros(128, K, dtype=torch.bfloat16, device=dev) if mode & 8 else None
    rc = lib.dsdf_selftest_umma(mode, A_dev.data_ptr(), B_dev.data_ptr(), Cout.data_ptr(),
                                store.data_ptr() if store is not None else None, K, N, a_lbo, a_sbo, b_lbo, b_sbo, st())
    if rc != 0:
        print(f"[selftest {tag}] launch error: {lib.dsdf_last_error().decode()}")
        return False
    try:
        torch.cuda.synchronize()
    except Exception as e:  # noqa: BLE001
        print(f"[selftest {tag}] CUDA error: {e}")
        return False
    err = (Cout.cpu() - ref).abs().max().item()
    ok = err < 1e-2 * max(1.0, ref.abs().max().item())
    extra = ""
    if store is not None:
        serr = (store.cpu().float() - A.float()).abs().max().item()
        extra = f" store_err={serr:.3g}"
        ok = ok and serr == 0.0
    print(f"[selftest {tag}] mode={mode} K={K} N={N} lbo/sbo A=({a_lbo},{a_sbo}) B=({b_lbo},{b_sbo}) max_err={err:.4g} "
          f"ref_max={ref.abs().max().item():.3g}{extra} -> {'OK' if ok else 'MISMATCH'}")
    return ok


def run_selftests():
    res = {}
    for K, N in ((64, 256), (128, 256), (256, 256), (256, 64), (128, 128)):
        res[f"kk_{K}_{N}"] = selftest(0, K, N, 16, 1024, 16, 1024, "K-major/K-major TMA")
    res["manualA"] = selftest(4, 256, 256, 16, 1024, 16, 1024, "manual sw128 A write")
    res["store"] = selftest(8, 256, 256, 16, 1024, 16, 1024, "TMA store of A tile")
    for K in (64, 128, 256):
        res[f"mn_{K}"] = selftest(3, K, 256, K * 128, 1024, K * 128, 1024, "MN-major A,B (LBO=group stride, SBO=1024)")
    if not all(res[f"mn_{K}"] for K in (64, 128, 256)):
        for K in (64, 128):
            selftest(3, K, 256, 1024, K * 128, 1024, K * 128, "MN-major alt (LBO=1024, SBO=group stride)")
    res["mn_a_only"] = selftest(1, 128, 256, 128 * 128, 1024, 16, 1024, "A MN-major, B K-major")
    res["mn_b_only"] = selftest(2, 128, 256, 16, 1024, 128 * 128, 1024, "A K-major, B MN-major")
    return res


def to_params_struct(spec, params):
    P = L.Params()
    keep = []
    for l in range(spec.num_linear):
        if spec.is_weight_normed(l):
            v = params[f"lin{l}.parametrizations.weight.original1"]
            g = params[f"lin{l}.parametrizations.weight.original0"]
            P.g[l] = g.data_ptr()
        else:
            v = params[f"lin{l}.weight"]
        P.v[l] = v.data_ptr()
        P.b[l] = params[f"lin{l}.bias"].data_ptr()
    return P


def fused_forward_check(name="big_8x512_L256", n_rows=1000, tag="eval", sdf_scale=True, timing=False):
    meta = json.load(open(os.path.join(ROOT, "tests", "golden", "decoder_cases.json")))[name]
    kw = meta["spec"]
    ospec = O.OracleSpec(**kw)
    spec = NetSpec(**kw)
    params_cpu = sdf_scale_params(ospec, meta["seed"]) if sdf_scale else O.make_params(ospec, meta["seed"])
    params = {k: v.to(dev).contiguous() for k, v in params_cpu.items()}
    desc = spec.desc()
    assert lib.dsdf_fused_supported(C.byref(desc)) == 1
    nbytes = lib.dsdf_pack_bytes(C.byref(desc))
    pack = torch.zeros(nbytes + 1024, dtype=torch.uint8, device=dev)
    off = (-pack.data_ptr()) % 1024
    pack = pack[off:off + nbytes]
    P = to_params_struct(spec, params)
    L.check(lib.dsdf_pack_weights(C.byref(desc), C.byref(P), pack.data_ptr(), st()))
    rng = np.random.RandomState(3)
    S_total, n_tiles = 5, (n_rows + 127) // 128
    table = torch.from_numpy((rng.normal(0, 1, size=(S_total, spec.latent_size)) / np.sqrt(spec.latent_size)).astype(np.float32))
    scene_ids = torch.tensor([3, 0, 4], dtype=torch.int64)
    tile_scene = torch.from_numpy(rng.randint(0, 3, size=n_tiles).astype(np.int32))
    xyz = torch.from_numpy(rng.uniform(-1, 1, size=(n_rows, 3)).astype(np.float32))
    table_d = table.to(dev)
    cbias = torch.zeros(3, 2, 512, device=dev)
    L.check(lib.dsdf_scene_prepare(C.byref(desc), pack.data_ptr(), table_d.data_ptr(), scene_ids.to(dev).data_ptr(), 3, 0.0,
                                   cbias.data_ptr(), st()))
    out = torch.full((n_rows,), float("nan"), device=dev)
    io = L.TrainIO()
    io.b_last = params[f"lin{spec.num_linear - 1}.bias"].data_ptr()
    io.training = 0
    xyz_d, ts_d = xyz.to(dev), tile_scene.to(dev)
    L.check(lib.dsdf_fused_forward(C.byref(desc), pack.data_ptr(), cbias.data_ptr(), ts_d.data_ptr(), xyz_d.data_ptr(), 3,
                                   n_rows, out.data_ptr(), C.byref(io), st()))
    try:
        torch.cuda.synchronize()
    except Exception as e:  # noqa: BLE001
        print(f"[fused fwd {name}] CUDA error: {e}")
        return False
    # oracle
    row_scene = scene_ids[tile_scene.long()].repeat_interleave(128)[:n_rows]
    inp = torch.cat([table[row_scene], xyz], 1)
    torch.set_num_threads(os.cpu_count())
    ref = O.decoder_forward(ospec, params_cpu, inp, training=False).squeeze(1)
    # scene-bias sanity (pure fp32 path)
    W0 = O.effective_weight(ospec, params_cpu, 0)
    c0_ref = table[scene_ids] @ W0[:, :spec.latent_size].t() + params_cpu["lin0.bias"]
    cerr = (cbias[:, 0, :].cpu() - c0_ref).abs().max().item()
    emu, _, _ = emulate_fused_forward(ospec, params_cpu, table, row_scene, xyz)
    emu_err = (out.cpu() - emu.squeeze(1)).abs().max().item()
    err = (out.cpu() - ref).abs()
    print(f"[fused fwd {name} {tag}] rows={n_rows} sdf_scale={sdf_scale} cbias_err={cerr:.3g} vs_bf16_emulation={emu_err:.3g} max_abs_err={err.max().item():.4g} mean_abs_err={err.mean().item():.4g} "
          f"ref range [{ref.min().item():.3f},{ref.max().item():.3f}] nan={int(torch.isnan(out).sum())} -> "
          f"{'OK' if err.max().item() < 2e-3 else 'MISMATCH'}")
    # timing of a larger eval run
    if timing:
        N = 1 << 20
        xyz_b = torch.rand(N, 3, device=dev) * 2 - 1
        ts_b = torch.zeros(N // 128, dtype=torch.int32, device=dev)
        out_b = torch.empty(N, device=dev)
        for _ in range(3):
            L.check(lib.dsdf_fused_forward(C.byref(desc), pack.data_ptr(), cbias.data_ptr(), ts_b.data_ptr(), xyz_b.data_ptr(), 3, N,
                                           out_b.data_ptr(), C.byref(io), st()))
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); e0.record()
        for _ in range(5):
            L.check(lib.dsdf_fused_forward(C.byref(desc), pack.data_ptr(), cbias.data_ptr(), ts_b.data_ptr(), xyz_b.data_ptr(), 3, N,
                                           out_b.data_ptr(), C.byref(io), st()))
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 5
        print(f"[fused fwd {name} timing] N={N} {ms:.3f} ms  -> {N / ms / 1e3:.2f} M rows/s, {N * 3.671e6 / ms / 1e9:.1f} algorithmic TFLOP/s")
    return err.max().item() < 2e-3


if __name__ == "__main__":
    t0 = time.time()
    print(torch.cuda.get_device_name(0), torch.cuda.get_device_capability(0))
    r = run_selftests()
    which = sys.argv[1:] or ["big_8x512_L256", "big_8x512_L2", "big_8x512_L16"]
    if r.get("kk_256_256") and r.get("manualA"):
        for nm in which:
            fused_forward_check(nm, sdf_scale=True, timing=(nm == which[0]))
            fused_forward_check(nm, sdf_scale=False)
    else:
        print("skipping fused forward: basic UMMA self-tests failed")
    print(f"done in {time.time() - t0:.1f}s")
