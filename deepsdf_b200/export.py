"""TorchScript export of a trained decoder: the job of ``create_libtorch_executable.py:4-23``.

The reference traces its ``Decoder`` on the CPU (``torch.jit.trace(decoder.eval().to("cpu"), example)``) and saves
``<experiment>/cpp_model.pt`` for its C++ consumers (libtorch, no Python, no CUDA library).  The drop-in ``Decoder`` here
computes in ``libdsdf_b200.so`` and has no CPU arithmetic to trace, so the export is written out explicitly: a graph-only
``nn.Module`` of aten ops restating ``deep_sdf_decoder.py:76-111`` in eval mode, with the weight-norm reparametrisation
folded once (``W = g * v / ||v||_row``, ``_weight_norm`` dim 0).  It is an EXCHANGE FORMAT for programs that cannot load
the CUDA library - nothing in this package ever calls it to compute (``Decoder.forward`` raises without the library), and
its output is checked against the reference's own ``cpp_model.pt`` outputs (``tests/golden/small_net_export.npz``).

    python -m deepsdf_b200.export -e <experiment_directory> [-c latest]      # writes <experiment_directory>/cpp_model.pt
"""
from __future__ import annotations

import os
from typing import Dict, Optional

import torch
from torch import nn


class ExportedDecoder(nn.Module):
    """Eval-mode graph of the decoder with folded weights.  Buffers ``weight{l}``, ``bias{l}`` (+ ``ln_w{l}``, ``ln_b{l}``)."""

    def __init__(self, spec, state: Dict[str, torch.Tensor]):
        super().__init__()
        self.num_linear = spec.num_linear
        self.geom_dimension = spec.geom_dimension
        self.latent_in = tuple(spec.latent_in or ())
        self.xyz_in_all = bool(spec.xyz_in_all)
        self.use_tanh = bool(spec.use_tanh)
        self.has_ln = []
        for l in range(spec.num_linear):
            pre = f"lin{l}."
            if pre + "weight" in state:
                w = state[pre + "weight"].detach().float()
            else:
                g = state[pre + "parametrizations.weight.original0"].detach().float()
                v = state[pre + "parametrizations.weight.original1"].detach().float()
                w = v * (g / v.norm(dim=1, keepdim=True))                       # torch._weight_norm(v, g, dim=0)
            self.register_buffer(f"weight{l}", w.clone())
            self.register_buffer(f"bias{l}", state[pre + "bias"].detach().float().clone())
            ln = f"bn{l}.weight" in state
            self.has_ln.append(ln)
            if ln:
                self.register_buffer(f"ln_w{l}", state[f"bn{l}.weight"].detach().float().clone())
                self.register_buffer(f"ln_b{l}", state[f"bn{l}.bias"].detach().float().clone())

    def forward(self, input: torch.Tensor) -> torch.Tensor:
        xyz = input[:, -self.geom_dimension:]
        x = input
        last = self.num_linear - 1
        for l in range(self.num_linear):
            if l in self.latent_in:
                x = torch.cat([x, input], 1)                                     # deep_sdf_decoder.py:88-89
            elif l != 0 and self.xyz_in_all:
                x = torch.cat([x, xyz], 1)                                       # :90-91
            x = torch.nn.functional.linear(x, getattr(self, f"weight{l}"), getattr(self, f"bias{l}"))
            if l == last:
                if self.use_tanh:
                    x = torch.tanh(x)                                            # :94-95
            else:
                if self.has_ln[l]:
                    w = getattr(self, f"ln_w{l}")
                    x = torch.nn.functional.layer_norm(x, (w.shape[0],), w, getattr(self, f"ln_b{l}"))   # :97-103
                x = torch.relu(x)                                                # :104 (dropout is the identity in eval mode)
        return torch.tanh(x)                                                     # :108-109 (always)


def exported_module(decoder) -> ExportedDecoder:
    """Graph-only module carrying ``decoder``'s current weights (CPU)."""
    state = {k: v.detach().cpu() for k, v in decoder.state_dict().items()}
    return ExportedDecoder(decoder.spec, state).eval()


def to_torchscript(decoder, example_input: Optional[torch.Tensor] = None, path: Optional[str] = None):
    """``torch.jit.trace`` + ``torch.jit.script`` of the exported graph like create_libtorch_executable.py:19-23; saves to
    ``path`` when given.  ``example_input`` [1, L+D] defaults to zeros."""
    m = exported_module(decoder)
    if example_input is None:
        example_input = torch.zeros(1, decoder.spec.latent_size + decoder.spec.geom_dimension)
    traced = torch.jit.trace(m, example_input.detach().cpu().float())
    sm = torch.jit.script(traced)
    if path is not None:
        sm.save(path)
    return sm


if __name__ == "__main__":
    import argparse
    from .reconstruct import load_decoder
    ap = argparse.ArgumentParser(description="Write <experiment>/cpp_model.pt (TorchScript of the trained decoder, eval mode)")
    ap.add_argument("--experiment_directory", "-e", required=True)
    ap.add_argument("--checkpoint", "-c", default="latest")
    a = ap.parse_args()
    dec, specs, epoch = load_decoder(a.experiment_directory, a.checkpoint, "cpu")
    lat_path = os.path.join(a.experiment_directory, "LatentCodes", a.checkpoint + ".pth")
    example = None
    if os.path.isfile(lat_path):                                                  # the reference traces with a trained code
        lat = torch.load(lat_path, map_location="cpu", weights_only=False)["latent_codes"]["weight"]
        example = torch.cat([lat[0].float(), torch.zeros(dec.spec.geom_dimension)]).unsqueeze(0)
    out = os.path.join(a.experiment_directory, "cpp_model.pt")
    to_torchscript(dec, example, out)
    print("wrote", out, "(epoch", epoch, ")")
