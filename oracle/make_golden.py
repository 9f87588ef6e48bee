#!/usr/bin/env python3
"""Generate tests/golden/*.npz from the REAL reference (runs only in the build container,
where /root/reference is mounted; the fixtures it writes are committed and travel).

  python oracle/make_golden.py [--reference /root/reference]

What is pinned:
  * small_net_export.npz  - the reference's own TorchScript export
    experiments/corner_spheres_only_small_network/cpp_model.pt (trained 4x64 / L=2 /
    latent_in=[1] / use_tanh network, created by create_libtorch_executable.py:19-23): its
    state_dict, 512 random inputs and the outputs the TorchScript graph itself produces.
  * decoder_cases.npz     - for a list of NetworkSpecs taken from experiments/*/specs.json
    (plus LayerNorm / xyz_in_all / latent_dropout variants), parameters from
    oracle.decoder_oracle.make_params(seed) loaded into the reference ``Decoder`` class
    (deep_sdf/networks/deep_sdf_decoder.py, imported by file path): eval outputs, the input
    gradient, and one training-step (train_deep_sdf.py:483-545 re-typed around the reference
    Decoder, nn.Embedding(max_norm), L1Loss(sum), torch.optim.Adam) with the decoder in
    eval() so that no RNG is involved: loss, latent grad, parameter-gradient norms/samples
    and the parameters after one Adam step.
"""
import argparse
import importlib.util
import json
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from oracle.decoder_oracle import OracleSpec, make_params  # noqa: E402

CASES = [
    # name, spec kwargs, n_rows
    ("small_4x64_L2_tanh", dict(latent_size=2, dims=[64] * 4, geom_dimension=3, dropout=list(range(8)), dropout_prob=0.2,
                                norm_layers=list(range(8)), latent_in=[1], weight_norm=True, xyz_in_all=False,
                                use_tanh=True, latent_dropout=False), 96),
    ("small_4x32_L1", dict(latent_size=1, dims=[32] * 4, geom_dimension=3, dropout=list(range(8)), dropout_prob=0.2,
                           norm_layers=list(range(8)), latent_in=[2], weight_norm=True, xyz_in_all=False,
                           use_tanh=False, latent_dropout=False), 96),
    ("mid_6x128_L1", dict(latent_size=1, dims=[128] * 6, geom_dimension=3, dropout=list(range(8)), dropout_prob=0.2,
                          norm_layers=list(range(8)), latent_in=[2], weight_norm=True, xyz_in_all=False,
                          use_tanh=False, latent_dropout=False), 64),
    ("big_8x512_L256", dict(latent_size=256, dims=[512] * 8, geom_dimension=3, dropout=list(range(8)), dropout_prob=0.2,
                            norm_layers=list(range(8)), latent_in=[4], weight_norm=True, xyz_in_all=False,
                            use_tanh=False, latent_dropout=False), 48),
    ("big_8x512_L2", dict(latent_size=2, dims=[512] * 8, geom_dimension=3, dropout=list(range(8)), dropout_prob=0.2,
                          norm_layers=list(range(8)), latent_in=[4], weight_norm=True, xyz_in_all=False,
                          use_tanh=False, latent_dropout=False), 48),
    ("big_8x512_L16", dict(latent_size=16, dims=[512] * 8, geom_dimension=3, dropout=list(range(8)), dropout_prob=0.2,
                           norm_layers=list(range(8)), latent_in=[4], weight_norm=True, xyz_in_all=False,
                           use_tanh=False, latent_dropout=False), 48),
    ("layernorm_4x48_L3", dict(latent_size=3, dims=[48] * 4, geom_dimension=3, dropout=[0, 1], dropout_prob=0.1,
                               norm_layers=[0, 1, 2, 3], latent_in=[2], weight_norm=False, xyz_in_all=False,
                               use_tanh=False, latent_dropout=False), 64),
    ("xyz_in_all_4x40_L4", dict(latent_size=4, dims=[40] * 4, geom_dimension=3, dropout=[0, 1, 2, 3], dropout_prob=0.2,
                                norm_layers=[0, 1, 2, 3], latent_in=[2], weight_norm=True, xyz_in_all=True,
                                use_tanh=True, latent_dropout=True), 64),
    ("geom2d_3x24_L2", dict(latent_size=2, dims=[24] * 3, geom_dimension=2, dropout=None, dropout_prob=0.0,
                            norm_layers=[], latent_in=[], weight_norm=False, xyz_in_all=False,
                            use_tanh=False, latent_dropout=False), 64),
]


def load_reference_decoder(ref_root):
    path = os.path.join(ref_root, "deep_sdf", "networks", "deep_sdf_decoder.py")
    spec = importlib.util.spec_from_file_location("ref_deep_sdf_decoder", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.Decoder


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reference", default="/root/reference")
    ap.add_argument("--out", default=os.path.join(os.path.dirname(HERE), "tests", "golden"))
    args = ap.parse_args()
    os.makedirs(args.out, exist_ok=True)
    torch.set_num_threads(1)  # bit-stable reductions
    RefDecoder = load_reference_decoder(args.reference)

    # ---- 1. the reference's own exported network --------------------------------------
    ts = torch.jit.load(os.path.join(args.reference, "experiments", "corner_spheres_only_small_network", "cpp_model.pt"),
                        map_location="cpu")
    ts.eval()
    sd = {k: v.detach().clone() for k, v in ts.state_dict().items()}
    rng = np.random.RandomState(1234)
    x = np.concatenate([rng.normal(0, 0.7, size=(512, 2)), rng.uniform(-1, 1, size=(512, 3))], 1).astype(np.float32)
    with torch.no_grad():
        y = ts(torch.from_numpy(x)).numpy()
    out = {"input": x, "output": y}
    for k, v in sd.items():
        out["param::" + k] = v.numpy()
    np.savez_compressed(os.path.join(args.out, "small_net_export.npz"), **out)
    print("small_net_export: keys", sorted(sd.keys()), "out range", y.min(), y.max())

    # ---- 2. spec cases through the reference Decoder class ------------------------------
    blob = {}
    meta = {}
    for ci, (name, kw, n_rows) in enumerate(CASES):
        spec = OracleSpec(**kw)
        seed = 100 + ci
        params = make_params(spec, seed)
        ref = RefDecoder(kw["latent_size"], **{k: v for k, v in kw.items() if k != "latent_size"})
        missing = ref.load_state_dict(params, strict=True)
        assert not missing.missing_keys and not missing.unexpected_keys
        ref.eval()
        rng = np.random.RandomState(seed + 7)
        L, D = spec.latent_size, spec.geom_dimension
        S_total, S = 6, 3
        sps = n_rows // S
        lat_scale = 1.0 / np.sqrt(L)
        table = (rng.normal(0, 1.0, size=(S_total, L)) * lat_scale).astype(np.float32)
        table[1] *= 3.0 / max(np.linalg.norm(table[1]), 1e-6)  # one row above CodeBound=1 -> exercises max_norm renorm
        scenes = np.array([4, 1, 2])
        scene_of_row = np.repeat(scenes, sps)
        xyz = rng.uniform(-1, 1, size=(S * sps, D)).astype(np.float32)
        gt = rng.normal(0, 0.08, size=(S * sps, 1)).astype(np.float32)

        # eval forward + input gradient on a dense [N, L+D] input
        inp = torch.from_numpy(np.concatenate([table[scene_of_row], xyz], 1)).requires_grad_(True)
        y = ref(inp)
        (G,) = torch.autograd.grad(y.sum(), inp)

        # one training step, decoder.eval() (no RNG), reference step math train_deep_sdf.py:483-545
        ref.zero_grad()
        lat_vecs = torch.nn.Embedding(S_total, L, max_norm=1.0)
        with torch.no_grad():
            lat_vecs.weight.copy_(torch.from_numpy(table))
        loss_l1 = torch.nn.L1Loss(reduction="sum")
        opt = torch.optim.Adam([{"params": ref.parameters(), "lr": 5e-4}, {"params": lat_vecs.parameters(), "lr": 1e-3}])
        clampd, lam, epoch = 0.1, 1e-4, 37
        sdf_gt = torch.clamp(torch.from_numpy(gt), -clampd, clampd)
        opt.zero_grad()
        batch_vecs = lat_vecs(torch.from_numpy(scene_of_row).long())
        input_t = torch.cat([batch_vecs, torch.from_numpy(xyz)], dim=1)
        pred = ref(input_t)
        pred_c = torch.clamp(pred, -clampd, clampd)
        n_total = xyz.shape[0]
        chunk_loss = loss_l1(pred_c, sdf_gt) / n_total
        l2_size_loss = torch.sum(torch.norm(batch_vecs, dim=1))
        chunk_loss = chunk_loss + (lam * min(1, epoch / 100) * l2_size_loss) / n_total
        chunk_loss.backward()
        grads = {k: p.grad.detach().clone() for k, p in ref.named_parameters()}
        lat_grad = lat_vecs.weight.grad.detach().clone()
        table_after_renorm = lat_vecs.weight.detach().clone()
        opt.step()
        post = {k: p.detach().clone() for k, p in ref.named_parameters()}
        post_lat = lat_vecs.weight.detach().clone()

        pre = f"{name}::"
        blob[pre + "table"] = table
        blob[pre + "scenes"] = scenes
        blob[pre + "xyz"] = xyz
        blob[pre + "gt"] = gt
        blob[pre + "eval_out"] = y.detach().numpy()
        blob[pre + "input_grad"] = G.numpy()
        blob[pre + "loss"] = np.array([float(chunk_loss.detach())], dtype=np.float64)
        blob[pre + "pred_train"] = pred.detach().numpy()
        blob[pre + "lat_grad"] = lat_grad.numpy()
        blob[pre + "table_after_renorm"] = table_after_renorm.numpy()
        blob[pre + "post_lat"] = post_lat.numpy()
        big = sum(int(np.prod(s)) for s in spec.param_shapes().values()) > 200_000
        srng = np.random.RandomState(5)
        for k in spec.param_names():
            g = grads[k].numpy()
            blob[pre + "gnorm::" + k] = np.array([np.linalg.norm(g.astype(np.float64))])
            if big:
                flat_idx = srng.randint(0, g.size, size=min(64, g.size))
                blob[pre + "gidx::" + k] = flat_idx
                blob[pre + "gsample::" + k] = g.reshape(-1)[flat_idx]
                blob[pre + "postsample::" + k] = post[k].numpy().reshape(-1)[flat_idx]
            else:
                blob[pre + "grad::" + k] = g
                blob[pre + "post::" + k] = post[k].numpy()
        meta[name] = dict(spec=kw, seed=seed, n_rows=int(S * sps), sps=int(sps), clamp=clampd, lam=lam, epoch=epoch,
                          lr=[5e-4, 1e-3], code_bound=1.0, big=bool(big))
        print(f"{name}: loss {float(chunk_loss):.6f} out[{y.min():.3f},{y.max():.3f}]")
    np.savez_compressed(os.path.join(args.out, "decoder_cases.npz"), **blob)
    with open(os.path.join(args.out, "decoder_cases.json"), "w") as f:
        json.dump(meta, f, indent=1)
    print("wrote", args.out)


if __name__ == "__main__":
    main()
