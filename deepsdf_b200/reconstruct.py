"""Batched frozen-decoder latent optimisation: upstream DeepSDF ``reconstruct.py`` (the fork keeps only its README entry,
``README.md:134-142``, and the helpers ``deep_sdf/data.py:66-71,113-139`` / ``deep_sdf/workspace.py:122-149``), run for
many shapes at once on the fused kernels.

Per iteration and shape (upstream defaults): draw ``num_samples`` rows of the shape's SDF samples, clamp prediction and
ground truth to +-clamp_dist, loss = mean |pred - gt| + 1e-4 * mean(z^2), Adam(lr 5e-3, x0.1 every num_iterations/2) on
the latent only; the decoder is in eval mode and frozen.  Shapes are independent, so a batch of S shapes is S
scene-aligned tile groups of one fused forward + dgrad pass; the latent gradient is the per-shape column sum of dY of
lin0 / the latent_in layer times the latent weight columns (``dsdf_fused_latent_finish``).  PARITY UNPINNED upstream
(no copy of reconstruct.py in the reference); checked against ``oracle.decoder_oracle.reconstruct``.
"""
from __future__ import annotations

from typing import List, Optional

import torch

from . import _lib as L
from .utils import fused_engine_for

TILE = 128


class LatentOptimizer:
    def __init__(self, decoder, num_shapes: int, num_samples: int = 8000, clamp_dist: float = 0.1, lr: float = 5e-3,
                 l2reg: bool = True, init_std: float = 0.01, seed: int = 0, init: Optional[torch.Tensor] = None):
        eng = fused_engine_for(decoder)
        if eng is None:
            raise L.DsdfError("LatentOptimizer needs a decoder supported by the fused path (8x512 class)")
        self.engine, self.device = eng, eng.device
        self.S, self.num_samples, self.clamp, self.lr0 = num_shapes, num_samples, clamp_dist, lr
        self.l2 = 1e-4 if l2reg else 0.0
        Lz = eng.spec.latent_size
        if init is None:
            g = torch.Generator().manual_seed(seed)
            init = torch.randn(num_shapes, Lz, generator=g) * init_std          # one N(0, init_std) code per shape
        self.latents = init.to(self.device, torch.float32).contiguous().clone()
        self.exp_avg = torch.zeros_like(self.latents)
        self.exp_avg_sq = torch.zeros_like(self.latents)
        self.grad = torch.zeros_like(self.latents)
        self.shape_loss = torch.zeros(num_shapes, device=self.device)
        self.tiles = (num_samples + TILE - 1) // TILE
        rows = self.tiles * TILE
        self.batch = torch.zeros(num_shapes, rows, 4, device=self.device)       # padded rows stay zero and are masked
        tr = torch.full((self.tiles,), TILE, dtype=torch.int32)
        tr[-1] = num_samples - (self.tiles - 1) * TILE
        self.tile_rows = tr.repeat(num_shapes).to(self.device)
        self.tile_scene = torch.arange(num_shapes, dtype=torch.int32, device=self.device).repeat_interleave(self.tiles)
        self.scene_ids = torch.arange(num_shapes, dtype=torch.int64, device=self.device)
        self.step_count = 0

    def step(self, samples: torch.Tensor, lr: float) -> torch.Tensor:
        """One Adam iteration on every shape.  ``samples`` [S, num_samples, 4] (x, y, z, sdf) device tensor: this
        iteration's subsample of each shape.  Returns the per-shape losses [S] (device tensor, pre-update)."""
        self.batch[:, :self.num_samples].copy_(samples)
        self.engine.latent_loss_and_grad(self.latents, self.scene_ids, self.tile_scene, self.tile_rows,
                                         self.batch.view(-1, 4), clamp=self.clamp, rows_per_shape=self.num_samples,
                                         l2reg=self.l2, lat_grad=self.grad, shape_loss=self.shape_loss)
        self.step_count += 1
        self.engine.adam(self.latents, self.grad, self.exp_avg, self.exp_avg_sq, self.step_count, lr)
        return self.shape_loss

    def reconstruct(self, data: torch.Tensor, num_iterations: int = 800, sample_idx: Optional[torch.Tensor] = None,
                    seed: int = 0, record_every: int = 0) -> List[torch.Tensor]:
        """``data`` [S, M, 4]: every shape's stored SDF samples (device).  ``sample_idx`` [iters, S, num_samples] fixes the
        subsamples (tests); otherwise they are drawn on the device.  Returns the recorded per-shape loss tensors."""
        data = data.to(self.device)
        gen = torch.Generator(device=self.device).manual_seed(seed)
        decreased_by, adjust_every = 10, int(num_iterations / 2)
        hist = []
        M = data.shape[1]
        for e in range(num_iterations):
            lr = self.lr0 * ((1 / decreased_by) ** (e // adjust_every))
            if sample_idx is not None:
                idx = sample_idx[e].to(self.device)
            else:
                idx = torch.randint(0, M, (self.S, self.num_samples), generator=gen, device=self.device)
            smp = torch.gather(data, 1, idx.unsqueeze(-1).expand(-1, -1, 4))
            loss = self.step(smp, lr)
            if record_every and (e % record_every == 0 or e == num_iterations - 1):
                hist.append(loss.clone())
        return hist
