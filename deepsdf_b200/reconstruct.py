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
        from .small import small_engine_for
        self.small = small_engine_for(decoder)            # 4x32 / 4x64 / 6x128: the fused fp32 small-network kernel
        eng = self.small if self.small is not None else fused_engine_for(decoder)
        if eng is None:
            raise L.DsdfError("LatentOptimizer needs a decoder of the fused 8x512 class or of the small-network class (hidden width <= 128)")
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
        from .small import adam_step
        if self.small is not None:
            self.small.latent_loss_and_grad(samples.contiguous(), self.latents, clamp=self.clamp, l2reg=self.l2, lat_grad=self.grad,
                                            shape_loss=self.shape_loss)
        else:
            self.batch[:, :self.num_samples].copy_(samples)
            self.engine.latent_loss_and_grad(self.latents, self.scene_ids, self.tile_scene, self.tile_rows,
                                             self.batch.view(-1, 4), clamp=self.clamp, rows_per_shape=self.num_samples,
                                             l2reg=self.l2, lat_grad=self.grad, shape_loss=self.shape_loss)
        self.step_count += 1
        adam_step(self.latents, self.grad, self.exp_avg, self.exp_avg_sq, self.step_count, lr)
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


# ---------------------------------------------------------------------------------------------- command line
reconstructions_subdir, reconstruction_codes_subdir, reconstruction_meshes_subdir = "Reconstructions", "Codes", "Meshes"


def load_decoder(experiment_directory: str, checkpoint: str, device):
    """deep_sdf/workspace.py:48-83 ``load_experiment_specifications`` + ``load_model_parameters``: returns (decoder in
    eval mode on ``device``, specs, saved epoch).  Checkpoints written under nn.DataParallel carry a ``module.`` prefix."""
    import json
    import os
    from .networks.deep_sdf_decoder import Decoder
    with open(os.path.join(experiment_directory, "specs.json")) as f:
        specs = json.load(f)
    decoder = Decoder(specs["CodeLength"], **specs["NetworkSpecs"])
    data = torch.load(os.path.join(experiment_directory, "ModelParameters", checkpoint + ".pth"), map_location="cpu", weights_only=False)
    sd = {(k[len("module."):] if k.startswith("module.") else k): v for k, v in data["model_state_dict"].items()}
    decoder.load_state_dict(sd)
    return decoder.to(device).eval(), specs, data["epoch"]


def reconstruct_split(experiment_directory: str, checkpoint: str = "latest", data_source: Optional[str] = None,
                      split_filename: Optional[str] = None, iters: int = 800, num_samples: int = 8000, skip: bool = False,
                      shapes_per_batch: int = 1024, resolution: int = 0, seed: int = 0, rank: int = 0, world: int = 1,
                      device="cuda") -> dict:
    """Upstream ``reconstruct.py -e <experiment> -c <checkpoint> -d <data> -s <split>``: optimise one latent code per
    instance of the split against the frozen decoder and write ``Reconstructions/<epoch>/Codes/<dataset>/<class>/
    <instance>.pth`` (a [1, 1, L] tensor like upstream's ``latent.unsqueeze(0)``; path helpers: deep_sdf/workspace.py:122-149).
    All shapes of a batch are optimised together; ranks take disjoint blocks of the instance list (no collective).
    Per-iteration samples come from the device store with the training sampler (half positive / half negative rows
    without replacement, deep_sdf/data.py:88-99) - upstream shuffles once and takes random contiguous windows
    (deep_sdf/data.py:113-139), the same distribution.  ``resolution`` > 0 also meshes every reconstruction like upstream
    (``create_mesh``: grid evaluation + iso-surface, both on the device) into ``Meshes/.../<instance>.ply``
    (deep_sdf/workspace.py:122-135)."""
    import json
    import os
    import numpy as np
    from . import sharding
    from .data import DeviceSDFSamples, sdf_samples_subdir
    from .mesh import create_mesh
    dev = torch.device(device)
    decoder, specs, saved_epoch = load_decoder(experiment_directory, checkpoint, dev)
    data_source = data_source or specs["DataSource"]
    split_filename = split_filename or specs.get("TestSplit") or specs["TrainSplit"]
    with open(split_filename) as f:
        split = json.load(f)
    instances = [(d, c, i) for d in split for c in split[d] for i in split[d][c]]
    mine = sharding.partition_scenes(len(instances), rank, world)
    clamp = specs.get("ClampingDistance", 0.1)
    root = os.path.join(experiment_directory, reconstructions_subdir, str(saved_epoch))
    todo = []
    for k in mine:
        d, c, i = instances[k]
        code_fn = os.path.join(root, reconstruction_codes_subdir, d, c, i + ".pth")
        if skip and os.path.isfile(code_fn):
            continue
        todo.append((d, c, i, code_fn))
    errs = {}
    for b0 in range(0, len(todo), shapes_per_batch):
        part = todo[b0:b0 + shapes_per_batch]
        arrays = []
        for d, c, i, _ in part:
            npz = np.load(os.path.join(data_source, sdf_samples_subdir, d, c, i + ".npz"))
            arrays.append((npz["pos"], npz["neg"]))
        store = DeviceSDFSamples(None, None, num_samples, 3, dev, arrays=arrays)
        S = len(part)
        opt = LatentOptimizer(decoder, S, num_samples=num_samples, clamp_dist=clamp, lr=5e-3, l2reg=True, init_std=0.01,
                              seed=seed + b0 + 7919 * rank)
        ids = torch.arange(S, dtype=torch.int64, device=dev)
        adjust_every, buf, loss = max(1, int(iters / 2)), None, None
        for e in range(iters):
            buf = store.sample_batch(ids, (seed * 0x9E3779B1 + e) & 0x7FFFFFFFFFFFFFFF, out=buf)
            loss = opt.step(buf.view(S, num_samples, 4), 5e-3 * (0.1 ** (e // adjust_every)))
        final = loss.cpu() if loss is not None else torch.zeros(S)
        lat = opt.latents.detach().cpu()
        for j, (d, c, i, code_fn) in enumerate(part):
            os.makedirs(os.path.dirname(code_fn), exist_ok=True)
            torch.save(lat[j].reshape(1, 1, -1).clone(), code_fn)
            errs[os.path.join(d, c, i)] = float(final[j])
            if resolution > 0:
                mesh_fn = os.path.join(root, reconstruction_meshes_subdir, d, c, i + ".ply")
                os.makedirs(os.path.dirname(mesh_fn), exist_ok=True)
                create_mesh(decoder, lat[j], mesh_fn, N=resolution)                      # writes the name as given (this fork's mesh.py:31)
    return {"epoch": saved_epoch, "loss": errs}


if __name__ == "__main__":
    import argparse
    import os
    ap = argparse.ArgumentParser(description="Use a trained DeepSDF decoder to reconstruct given SDF samples (fused B200 path)")
    ap.add_argument("--experiment", "-e", dest="experiment_directory", required=True)
    ap.add_argument("--checkpoint", "-c", default="latest")
    ap.add_argument("--data", "-d", dest="data_source", default=None)
    ap.add_argument("--split", "-s", dest="split_filename", default=None)
    ap.add_argument("--iters", type=int, default=800)
    ap.add_argument("--num-samples", type=int, default=8000)
    ap.add_argument("--skip", action="store_true", help="skip instances whose code file already exists")
    ap.add_argument("--shapes-per-batch", type=int, default=1024)
    ap.add_argument("--resolution", type=int, default=0, help="also mesh every reconstruction on an N^3 grid (Meshes/.../<instance>.ply)")
    ap.add_argument("--seed", type=int, default=0)
    a = ap.parse_args()
    if not torch.cuda.is_available():
        raise SystemExit("deepsdf_b200.reconstruct needs a CUDA device (sm_100a); there is no CPU path")
    rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", "0"), ("WORLD_SIZE", "1"), ("LOCAL_RANK", "0")))
    torch.cuda.set_device(local)
    res = reconstruct_split(a.experiment_directory, a.checkpoint, a.data_source, a.split_filename, a.iters, a.num_samples,
                            a.skip, a.shapes_per_batch, a.resolution, a.seed, rank, world, torch.device("cuda", local))
    print("rank {}: reconstructed {} shapes of epoch {}, mean loss {:.5f}".format(
        rank, len(res["loss"]), res["epoch"], sum(res["loss"].values()) / max(1, len(res["loss"]))))
