"""NetworkSpecs -> resolved layer table and the C-ABI ``dsdf_net_desc``.

Mirrors what ``Decoder.__init__`` derives from its keyword arguments
(reference deep_sdf/networks/deep_sdf_decoder.py:10-73): layer in/out sizes with the
``latent_in`` narrowing, which layers are weight-normed / LayerNormed / followed by dropout.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Optional, Sequence

from ._lib import MAX_LAYERS, NetDesc


@dataclass
class NetSpec:
    latent_size: int
    dims: Sequence[int]
    geom_dimension: int
    dropout: Optional[Sequence[int]] = None
    dropout_prob: float = 0.0
    norm_layers: Sequence[int] = ()
    latent_in: Sequence[int] = ()
    weight_norm: bool = False
    xyz_in_all: Optional[bool] = None
    use_tanh: bool = False
    latent_dropout: bool = False

    def __post_init__(self):
        full = [self.latent_size + self.geom_dimension] + list(self.dims) + [1]   # decoder.py:29
        self.full_dims: List[int] = full
        self.num_layers = len(full)                                               # decoder.py:31 (reference attribute)
        n = len(full)
        self.in_dims: List[int] = []
        self.out_dims: List[int] = []
        for layer in range(n - 1):                                                # decoder.py:42-48
            if layer + 1 in (self.latent_in or ()):
                out_dim = full[layer + 1] - full[0]
            else:
                out_dim = full[layer + 1]
                if self.xyz_in_all and layer != n - 2:
                    out_dim -= self.geom_dimension
            self.in_dims.append(full[layer])
            self.out_dims.append(out_dim)
            if out_dim <= 0:
                raise ValueError(f"layer {layer} would have {out_dim} outputs")
        if self.num_linear > MAX_LAYERS:
            raise ValueError(f"at most {MAX_LAYERS} linear layers are supported")

    @property
    def num_linear(self) -> int:
        return len(self.full_dims) - 1

    def is_weight_normed(self, layer: int) -> bool:
        return bool(self.weight_norm) and layer in (self.norm_layers or ())        # decoder.py:50

    def has_layernorm(self, layer: int) -> bool:
        return (not self.weight_norm) and self.norm_layers is not None and layer in self.norm_layers   # decoder.py:59-64

    def layernorm_applied(self, layer: int) -> bool:
        return self.has_layernorm(layer) and layer < self.num_linear - 1           # decoder.py:96-103

    def has_dropout(self, layer: int) -> bool:
        return (self.dropout is not None and layer in self.dropout and layer < self.num_linear - 1
                and self.dropout_prob > 0.0)                                       # decoder.py:105-106

    def uses_dropout(self) -> bool:
        return any(self.has_dropout(layer) for layer in range(self.num_linear - 1))

    def desc(self) -> NetDesc:
        d = NetDesc()
        d.latent_size = self.latent_size
        d.geom_dim = self.geom_dimension
        d.num_linear = self.num_linear
        wn = ln = dr = li = 0
        for l in range(self.num_linear):
            d.in_dim[l] = self.in_dims[l]
            d.out_dim[l] = self.out_dims[l]
            wn |= int(self.is_weight_normed(l)) << l
            ln |= int(self.layernorm_applied(l)) << l
            dr |= int(self.has_dropout(l)) << l
            li |= int(l in (self.latent_in or ())) << l
        d.weight_norm_mask, d.layernorm_mask, d.dropout_mask, d.latent_in_mask = wn, ln, dr, li
        d.xyz_in_all = int(bool(self.xyz_in_all))
        d.use_tanh = int(bool(self.use_tanh))
        d.latent_dropout = int(bool(self.latent_dropout))
        d.dropout_prob = float(self.dropout_prob)
        return d
