"""Shared helpers for the parity tests."""
import json
import os

import numpy as np
import torch

from oracle import decoder_oracle as O

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def load_cases():
    with open(os.path.join(GOLDEN, "decoder_cases.json")) as f:
        meta = json.load(f)
    blob = np.load(os.path.join(GOLDEN, "decoder_cases.npz"))
    return meta, blob


def build_module(kw, params=None, device=None):
    from deepsdf_b200.networks.deep_sdf_decoder import Decoder
    m = Decoder(kw["latent_size"], **{k: v for k, v in kw.items() if k != "latent_size"})
    if params is not None:
        res = m.load_state_dict(params, strict=True)
        assert not res.missing_keys and not res.unexpected_keys
    if device is not None:
        m = m.to(device)
    return m


def build_exact_module(kw, params=None, device=None):
    """The drop-in module pinned to its exact fp32 mode: eval-mode forwards of fused-class networks with CodeLength <= 16 stay
    on the generic CUDA-core path instead of the bf16 tensor-core chain (``Decoder.fused_eval``)."""
    m = build_module(kw, params, device)
    m.fused_eval = False
    return m


def sdf_scale_params(ospec, seed, out_scale=0.08):
    """Synthetic parameters whose outputs live on the SDF scale (|sdf| <~ 0.2), like a trained DeepSDF network:
    He-gain hidden layers and a last layer shrunk so that the pre-tanh output has std ~ out_scale."""
    p = O.make_params(ospec, seed)
    n = ospec.num_linear - 1
    key = f"lin{n}.parametrizations.weight.original0" if ospec.is_weight_normed(n) else f"lin{n}.weight"
    rng = np.random.RandomState(seed + 999)
    L, D = ospec.latent_size, ospec.geom_dimension
    x = torch.from_numpy(np.concatenate([rng.normal(0, 1 / np.sqrt(L), (256, L)), rng.uniform(-1, 1, (256, D))], 1).astype(np.float32))
    q = {k: v.clone() for k, v in p.items()}
    # measure the pre-tanh std with the last layer at its synthetic value
    y = O.decoder_forward(ospec, q, x)
    pre = torch.atanh(y.clamp(-0.999999, 0.999999))
    if ospec.use_tanh:
        pre = torch.atanh(pre.clamp(-0.999999, 0.999999))
    s = float(pre.std()) + 1e-6
    q[key] = q[key] * (out_scale / s)
    q[f"lin{n}.bias"] = q[f"lin{n}.bias"] * 0.0 + 0.01
    return q


def bf16r(x):
    return x.to(torch.bfloat16).float()


def emulate_fused_forward(ospec, params, table, row_scene, xyz, masks=None):
    """CPU emulation of the fused tcgen05 forward's arithmetic: bf16 weights and bf16 inter-layer activations with fp32
    accumulation for the hidden-activation columns; latent and xyz columns, biases and the last layer in fp32.
    Returns (sdf[N,1], pre-tanh y[N,1], list of post-activation fp32 tensors before bf16 rounding)."""
    L, D = ospec.latent_size, ospec.geom_dimension
    nh = ospec.num_linear - 1
    z = table[row_scene]
    a = None
    acts = []
    pend = 1.0      # dropout scale of the previous layer: folded into this layer's activation columns before their bf16 rounding
    for l in range(nh):
        W = O.effective_weight(ospec, params, l)
        b = params[f"lin{l}.bias"]
        if l == 0:
            h = z @ W[:, :L].t() + b + xyz @ W[:, L:L + D].t()
        elif l in ospec.latent_in:
            xw = ospec.in_dims[l] - (L + D)
            h = a[:, :xw] @ bf16r(W[:, :xw] * pend).t() + (z @ W[:, xw:xw + L].t() + b) + xyz @ W[:, xw + L:].t()
        else:
            h = a @ bf16r(W * pend).t() + b
        h = torch.relu(h)
        pend = 1.0
        if masks is not None and l in masks:
            h = h * masks[l]
            pend = ospec.drop_scale()
        acts.append(h * pend)
        a = bf16r(h)
    Wl = O.effective_weight(ospec, params, nh)
    y = acts[-1] @ Wl.t() + params[f"lin{nh}.bias"]
    t = torch.tanh(y)
    if ospec.use_tanh:
        t = torch.tanh(t)
    return t, y, acts


def unpack_gate_bits(bitmask: torch.Tensor, ospec, n_rows: int):
    """The fused forward's 1-bit relu*dropout gates (dsdf_train_io.bitmask: uint32 [num_hidden][n_tiles][16 words][128 rows],
    word w bit i = column 32*w + i) -> {layer: float {0,1} [n_rows, out_dim]} on the CPU."""
    nh = ospec.num_linear - 1
    n_tiles = (n_rows + 127) // 128
    words = bitmask.view(torch.int32)[:nh * n_tiles * 16 * 128].reshape(nh, n_tiles, 16, 128).permute(0, 1, 3, 2)
    words = words.reshape(nh, n_tiles * 128, 16)[:, :n_rows].cpu().numpy().astype(np.uint32)
    out = {}
    for l in range(nh):
        bits = ((words[l][:, :, None] >> np.arange(32, dtype=np.uint32)[None, None, :]) & 1).reshape(n_rows, 512)
        out[l] = torch.from_numpy(bits[:, :ospec.out_dims[l]].astype(np.float32))
    return out
