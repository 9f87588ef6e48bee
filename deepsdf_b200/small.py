"""Host-side driver of the small-network kernel (``csrc/smallnet.cu``): the 4x32 / 4x64 / 6x128 decoders the reference's
optimisation and analysis consumers run (``optimization_runs/*/config.json:8`` -> ``experiments/double_lattice_3D_small_network``,
``round_cross_big_network`` ...).  One launch evaluates the SDF of a batch of queries and, on request, the input gradient
``G = dSDF/d[latent, xyz]`` of every row - the quantity ``deep_sdf/mesh.py:405-422`` re-derives with L x n_cpt ``jvp`` calls.

fp32 throughout (FFMA, fixed summation order): results agree with the reference's fp32 arithmetic to ~1e-6.  Eval mode only
(the consumers call ``decoder.eval()``); training of these networks runs on the generic path (``Decoder.forward``).
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Tuple

import torch

from . import _lib as L


def param_key(decoder) -> tuple:
    """Identity + version of every parameter tensor (torch bumps ``_version`` on in-place updates; kernels that write the
    parameters behind torch's back bump ``decoder._param_epoch``): the cache key of a weight pack."""
    return (tuple((p.data_ptr(), p._version) for p in decoder.parameters()), getattr(decoder, "_param_epoch", 0))


class SmallEngine:
    def __init__(self, decoder):
        self.lib = L.lib()
        self.decoder, self.spec, self.desc = decoder, decoder.spec, decoder._desc
        if self.lib.dsdf_small_supported(C.byref(self.desc)) != 1:
            raise L.DsdfError("this NetworkSpecs is not supported by the small-network kernel (hidden widths <= 128, no LayerNorm / "
                              "xyz_in_all / latent_dropout)")
        p0 = next(decoder.parameters())
        L.require_cuda(p0)
        self.device = p0.device
        with torch.cuda.device(self.device):
            L.check(self.lib.dsdf_check_device())
        self.pack = torch.zeros(self.lib.dsdf_small_pack_bytes(C.byref(self.desc)) // 4, device=self.device)
        self._key = None
        self.repack()

    def repack(self) -> None:
        """Fold g*v/||v|| and rebuild the padded fp32 operand blocks - only when the parameters changed since the last pack."""
        key = param_key(self.decoder)
        if key == self._key:
            return
        P = self.decoder.c_params()
        with torch.cuda.device(self.device):
            L.check(self.lib.dsdf_small_pack_weights(C.byref(self.desc), C.byref(P), self.pack.data_ptr(), L.stream_ptr(self.device)))
        self._key = key

    def sdf_and_input_grad(self, queries: torch.Tensor, latent: Optional[torch.Tensor] = None, with_grad: bool = True,
                           sdf: Optional[torch.Tensor] = None, G: Optional[torch.Tensor] = None
                           ) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
        """Eval-mode SDF [N] and (``with_grad``) G [N, L+D] = dSDF/d[latent, xyz] (column order of the decoder input).
        ``queries`` [N, L+D] rows carry their own code; with ``latent`` [L] they are [N, D] points and the one code is
        broadcast (``decode_sdf(decoder, latent_vector, queries)``, deep_sdf/utils.py:57-61)."""
        L.require_cuda(queries)
        Lz, D = self.spec.latent_size, self.spec.geom_dimension
        N, width = queries.shape
        if queries.dtype != torch.float32 or queries.stride(1) != 1:
            raise L.DsdfError("queries must be fp32 with unit column stride")
        if width != (D if latent is not None else Lz + D):
            raise L.DsdfError(f"queries must be [N, {D if latent is not None else Lz + D}]")
        sdf = sdf if sdf is not None else torch.empty(N, device=self.device)
        if with_grad:
            G = G if G is not None else torch.empty(N, Lz + D, device=self.device)
            assert G.dtype == torch.float32 and G.shape[0] >= N and G.shape[1] >= Lz + D and G.stride(1) == 1
        else:
            G = None
        if N == 0:
            return sdf, G
        lat = None
        if latent is not None:
            lat = latent.detach().reshape(-1).to(self.device, torch.float32).contiguous()
            if lat.numel() != Lz:
                raise L.DsdfError(f"latent must have {Lz} entries")
        self.repack()
        with torch.cuda.device(self.device):
            L.check(self.lib.dsdf_small_forward(C.byref(self.desc), self.pack.data_ptr(), queries.data_ptr(), queries.stride(0),
                                                lat.data_ptr() if lat is not None else None, N, sdf.data_ptr(),
                                                G.data_ptr() if G is not None else None, G.stride(0) if G is not None else 0,
                                                L.stream_ptr(self.device)))
        return sdf, G

    def latent_loss_and_grad(self, samples: torch.Tensor, latents: torch.Tensor, *, clamp: float, l2reg: float,
                             lat_grad: torch.Tensor, shape_loss: torch.Tensor) -> None:
        """One iteration of upstream reconstruct.py for every shape: ``samples`` [S, R, 4] (x, y, z, sdf) device tensor,
        ``latents`` [S, L].  Writes lat_grad [S, L] = d(mean|clamp(pred)-clamp(gt)| + l2reg*mean(z^2))/dz and shape_loss [S]."""
        L.require_cuda(samples, latents)
        S, R, w = samples.shape
        Lz = self.spec.latent_size
        assert w == 4 and samples.is_contiguous() and samples.dtype == torch.float32 and latents.shape == (S, Lz)
        key = (S, R)
        if getattr(self, "_lat_key", None) != key:
            self._lat_scratch = (torch.empty(S * R, device=self.device), torch.empty(S * R, Lz + 3, device=self.device),
                                 torch.empty(S * R, device=self.device))
            self._lat_key = key
        sdf, G, rl = self._lat_scratch
        self.repack()
        with torch.cuda.device(self.device):
            L.check(self.lib.dsdf_small_latent_step(C.byref(self.desc), self.pack.data_ptr(), samples.data_ptr(), 4, latents.data_ptr(),
                                                    S, R, float(clamp), float(l2reg), sdf.data_ptr(), G.data_ptr(), rl.data_ptr(),
                                                    lat_grad.data_ptr(), shape_loss.data_ptr(), L.stream_ptr(self.device)))


def adam_step(p: torch.Tensor, g: torch.Tensor, m: torch.Tensor, v: torch.Tensor, step: int, lr: float,
              betas=(0.9, 0.999), eps: float = 1e-8) -> None:
    """torch.optim.Adam update of one flat fp32 tensor through ``dsdf_adam`` (shared by the engines)."""
    with torch.cuda.device(p.device):
        L.check(L.lib().dsdf_adam(p.data_ptr(), g.data_ptr(), m.data_ptr(), v.data_ptr(), p.numel(), float(lr), betas[0], betas[1],
                                  eps, int(step), 1.0, None, L.stream_ptr(p.device)))


def small_engine_for(decoder) -> Optional[SmallEngine]:
    """The module's cached SmallEngine, or None if its NetworkSpecs is outside the small-network kernel's class."""
    dec = decoder.module if isinstance(decoder, torch.nn.DataParallel) else decoder
    if not hasattr(dec, "_desc") or L.lib().dsdf_small_supported(C.byref(dec._desc)) != 1:
        return None
    eng = getattr(dec, "_small_engine", None)
    if eng is None or eng.device != next(dec.parameters()).device:
        eng = SmallEngine(dec)
        object.__setattr__(dec, "_small_engine", eng)
    else:
        eng.repack()
    return eng
