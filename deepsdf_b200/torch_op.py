"""``torch.library`` registration of the decoder's eval-mode forward, so that ``torch.jit.trace(decoder, example_input)``
(``create_libtorch_executable.py:19-20``) records ONE graph node, ``deepsdf_b200::decoder_eval``, instead of failing on an
``autograd.Function`` over ctypes.  The node calls the same C-ABI entry point as ``Decoder.forward`` (generic fp32 path:
any NetworkSpecs), so a traced / scripted module reproduces the module's outputs bit for bit wherever this package is
imported (the op lives in this process; a plain-libtorch consumer without the library uses ``deepsdf_b200.export``'s
explicit aten graph instead).

The network description travels as plain ints (schema types), the parameters as a tensor list in ``flat_param_list()`` order.
"""
from __future__ import annotations

import ctypes as C
from typing import List

import torch

from . import _lib as L
from .spec import NetSpec

_NAME = "deepsdf_b200::decoder_eval"


def encode_spec(s: NetSpec) -> List[int]:
    def mask(xs):
        return sum(1 << int(i) for i in (xs or ()) if 0 <= int(i) < 32)
    return [s.latent_size, s.geom_dimension, len(s.dims)] + [int(d) for d in s.dims] + [
        mask(s.dropout) if s.dropout is not None else -1, mask(s.norm_layers), mask(s.latent_in), int(bool(s.weight_norm)),
        int(bool(s.xyz_in_all)), int(bool(s.use_tanh)), int(bool(s.latent_dropout))]


def decode_spec(code: List[int], dropout_prob: float) -> NetSpec:
    L_, D, n = code[0], code[1], code[2]
    dims = list(code[3:3 + n])
    dm, nm, lm, wn, xa, ut, ld = code[3 + n:3 + n + 7]
    bits = lambda m: [i for i in range(32) if (m >> i) & 1]      # noqa: E731
    return NetSpec(L_, dims, D, None if dm < 0 else bits(dm), dropout_prob, bits(nm), bits(lm), bool(wn), bool(xa), bool(ut), bool(ld))


@torch.library.custom_op(_NAME, mutates_args=())
def decoder_eval(inp: torch.Tensor, params: List[torch.Tensor], spec_code: List[int], dropout_prob: float) -> torch.Tensor:
    spec = decode_spec(list(spec_code), dropout_prob)
    desc = spec.desc()
    L.require_cuda(inp, *params)
    lib = L.lib()
    x = inp.detach().contiguous().float()
    N = x.shape[0]
    out = torch.empty(N, 1, device=x.device, dtype=torch.float32)
    if N == 0:
        return out
    P = L.Params()
    it = iter(params)
    for layer in range(spec.num_linear):
        names = ["v"] + (["g"] if spec.is_weight_normed(layer) else []) + ["b"] + (["ln_w", "ln_b"] if spec.has_layernorm(layer) else [])
        for nm in names:
            t = next(it)
            if t.dtype != torch.float32 or not t.is_contiguous():
                raise L.DsdfError("decoder parameters must be contiguous fp32")
            getattr(P, nm)[layer] = t.data_ptr()
    with torch.cuda.device(x.device):
        nbytes = lib.dsdf_generic_workspace_bytes(C.byref(desc), N)
        ws = torch.empty(nbytes, dtype=torch.uint8, device=x.device)
        L.check(lib.dsdf_generic_forward(C.byref(desc), C.byref(P), x.data_ptr(), N, 0, 0, 0, out.data_ptr(), ws.data_ptr(), nbytes,
                                         L.stream_ptr(x.device)))
    return out


@decoder_eval.register_fake
def _(inp, params, spec_code, dropout_prob):
    return inp.new_empty((inp.shape[0], 1), dtype=torch.float32)
