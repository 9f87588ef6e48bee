#!/usr/bin/env python3
"""Thin multi-GPU training entry point on the fused path: ``train_deep_sdf.py -e <experiment_directory>`` semantics
(train_deep_sdf.py:255-581) with the reference's experiment layout and checkpoint formats, one process per GPU.

  python -m deepsdf_b200.train -e experiments/foo                       # one GPU
  python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 -m deepsdf_b200.train -e experiments/foo

Reads ``specs.json`` (DataSource, TrainSplit, NetworkSpecs, CodeLength, NumEpochs, SnapshotFrequency, AdditionalSnapshots,
LearningRateSchedule, SamplesPerScene, ScenesPerBatch, ClampingDistance, CodeRegularization(+Lambda), CodeBound,
CodeInitStdDev, GradientClipNorm, LogFrequency).  Scenes are block-partitioned over the ranks (``sharding``); every rank
loads ITS scenes' npz files into HBM once (``DeviceSDFSamples``) and steps ``ScenesPerBatch / world_size`` scenes per
iteration; the decoder gradient is all-reduced, latent codes never move until a checkpoint gathers them.
``--strict-shuffle`` (or ``"SceneSharding": "replicated"`` in specs.json) is the strict mode of SURVEY.md 8(e): every rank
holds all scenes and all latent rows, batches come from the reference's global shuffle, rank r steps the r-th share of
each batch and the latent gradient is all-reduced with the decoder gradient - the N-GPU run then reproduces the 1-GPU
loss curve up to summation order (tests/test_multirank_gpu.py).

Rank 0 writes what the reference writes, loadable by the reference's own tools (``deep_sdf/workspace.py``):
  ModelParameters/<epoch|latest>.pth   {"epoch", "model_state_dict"} with DataParallel's ``module.`` key prefix  (:96-104)
  LatentCodes/<epoch|latest>.pth       {"epoch", "latent_codes": {"weight": [num_scenes, L]}}                    (:134-143)
  OptimizerParameters/<...>.pth        {"epoch", "optimizer_state_dict": flat Adam moments of this trainer}      (:106-113)
  Logs.pth                             {"epoch", "loss", "learning_rate", "timing", "latent_magnitude", "param_magnitude"}
Python here is plumbing only; all arithmetic is in libdsdf_b200.so."""
from __future__ import annotations

import argparse
import json
import logging
import os
import time

import torch

from . import sharding
from .data import DeviceSDFSamples
from .trainer import SceneTrainer, get_spec_with_default

model_params_subdir, optimizer_params_subdir, latent_codes_subdir = "ModelParameters", "OptimizerParameters", "LatentCodes"
logs_filename, specifications_filename = "Logs.pth", "specs.json"             # deep_sdf/workspace.py:9-16


def split_subset(split: dict, lo: int, hi: int) -> dict:
    """Instances [lo, hi) of a split file, in the order ``get_instance_filenames`` enumerates them (data.py:15-33)."""
    out, i = {}, 0
    for dataset in split:
        for class_name in split[dataset]:
            for inst in split[dataset][class_name]:
                if lo <= i < hi:
                    out.setdefault(dataset, {}).setdefault(class_name, []).append(inst)
                i += 1
    return out


def count_instances(split: dict) -> int:
    return sum(len(split[d][c]) for d in split for c in split[d])


def gather_latents(trainer: SceneTrainer, num_scenes: int, world: int, rank: int) -> torch.Tensor:
    """All ranks' latent rows in global scene order (rank r owns the block ``partition_scenes(num_scenes, r, world)``)."""
    if world == 1 or trainer.replicated:
        return trainer.lat_vecs.detach().cpu().clone()
    import torch.distributed as dist
    most = max(len(sharding.partition_scenes(num_scenes, r, world)) for r in range(world))
    pad = torch.zeros(most, trainer.latent_size, device=trainer.device)
    pad[:trainer.lat_vecs.shape[0]] = trainer.lat_vecs
    parts = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(parts, pad)
    return torch.cat([p[:len(sharding.partition_scenes(num_scenes, r, world))] for r, p in enumerate(parts)], 0).cpu()


def save_checkpoint(exp: str, name: str, trainer: SceneTrainer, all_latents: torch.Tensor, epoch: int) -> None:
    for sub in (model_params_subdir, optimizer_params_subdir, latent_codes_subdir):
        os.makedirs(os.path.join(exp, sub), exist_ok=True)
    sd = {"module." + k: v.detach().cpu().clone() for k, v in trainer.decoder.state_dict().items()}
    torch.save({"epoch": epoch, "model_state_dict": sd}, os.path.join(exp, model_params_subdir, name))
    torch.save({"epoch": epoch, "latent_codes": {"weight": all_latents}}, os.path.join(exp, latent_codes_subdir, name))
    torch.save({"epoch": epoch, "optimizer_state_dict": {
        "format": "deepsdf_b200 flat Adam", "step": trainer.step_count, "param_names": trainer.decoder.flat_param_names(),
        "offsets": trainer._offs, "exp_avg": trainer.exp_avg.cpu(), "exp_avg_sq": trainer.exp_avg_sq.cpu(),
        "lat_exp_avg": trainer.lat_exp_avg.cpu(), "lat_exp_avg_sq": trainer.lat_exp_avg_sq.cpu()}},
        os.path.join(exp, optimizer_params_subdir, name))


def main_function(experiment_directory: str, data_source: str = None, max_epochs: int = None,
                  strict_shuffle: bool = None, backend: str = "nccl") -> dict:
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("deepsdf_b200.train needs a CUDA device (sm_100a); there is no CPU path")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        if not dist.is_initialized():
            dist.init_process_group(backend, **({"device_id": dev} if backend == "nccl" else {}))
    logging.basicConfig(level=logging.INFO if rank == 0 else logging.WARNING, format="%(asctime)s %(message)s", datefmt="%H:%M:%S")
    with open(os.path.join(experiment_directory, specifications_filename)) as f:
        specs = json.load(f)
    logging.info("Experiment description: \n" + str(specs.get("Description", "")))
    data_source = data_source or specs["DataSource"]
    with open(specs["TrainSplit"]) as f:
        train_split = json.load(f)
    num_scenes = count_instances(train_split)
    if strict_shuffle is None:
        strict_shuffle = get_spec_with_default(specs, "SceneSharding", "partition") == "replicated"
    mine = range(num_scenes) if strict_shuffle else sharding.partition_scenes(num_scenes, rank, world)
    sps, spb = specs["SamplesPerScene"], specs["ScenesPerBatch"]
    if spb % world:
        raise SystemExit(f"ScenesPerBatch ({spb}) must be a multiple of the number of GPUs ({world})")
    spb_local = spb // world
    dataset = DeviceSDFSamples(data_source, split_subset(train_split, mine.start, mine.stop), sps, 3, dev)
    logging.info("There are {} scenes ({} on this GPU, {} rows resident)".format(num_scenes, len(dataset), dataset.total_rows))
    trainer = SceneTrainer(specs, len(dataset), dev, rank=rank, world_size=world, seed=get_spec_with_default(specs, "Seed", 0),
                           replicated=strict_shuffle)
    num_epochs = specs["NumEpochs"] if max_epochs is None else min(max_epochs, specs["NumEpochs"])
    checkpoints = sorted(set(list(range(specs["SnapshotFrequency"], specs["NumEpochs"] + 1, specs["SnapshotFrequency"]))
                             + list(specs.get("AdditionalSnapshots", []))))                     # :281-291
    log_frequency = get_spec_with_default(specs, "LogFrequency", 10)
    loss_log, lr_log, timing_log, lat_mag_log, param_mag_log = [], [], [], [], {}
    gen = torch.Generator().manual_seed(1234 + (0 if strict_shuffle else rank))
    its = num_scenes // spb if strict_shuffle else len(dataset) // spb_local
    if world > 1 and not strict_shuffle:               # every rank must take the same number of iterations per epoch
        import torch.distributed as dist
        t = torch.tensor([its], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        its = int(t.item())
    for epoch in range(1, num_epochs + 1):
        start = time.time()
        if strict_shuffle:
            losses = trainer.train_epoch(dataset, epoch, scenes_per_batch=spb, generator=gen, max_iterations=its)
        else:
            losses = trainer.train_epoch(dataset, epoch, scenes_per_batch=spb_local, generator=gen, max_iterations=its,
                                         global_rows=spb * sps, row0=sharding.row_offset(rank, spb_local * sps))
        ep = torch.stack([l[0] for l in losses]) if losses else torch.zeros(0, device=dev)
        if world > 1:
            import torch.distributed as dist
            dist.all_reduce(ep)                                                                  # ranks hold partial sums of the batch loss
        loss_log += [float(x) for x in ep.cpu()]                                                 # one sync per epoch, not per chunk (:535)
        timing_log.append(time.time() - start)
        lr_log.append(trainer.learning_rates(epoch))
        all_lat = None
        if epoch in checkpoints or epoch % log_frequency == 0 or epoch == num_epochs:
            all_lat = gather_latents(trainer, num_scenes, world, rank)
        if rank == 0:
            lat_mag_log.append(trainer.mean_latent_magnitude())
            for n, p in trainer.decoder.named_parameters():                                      # :246-252
                param_mag_log.setdefault(n[7:] if n.startswith("module.") else n, []).append(float(p.data.norm()))
            if epoch in checkpoints:
                save_checkpoint(experiment_directory, str(epoch) + ".pth", trainer, all_lat, epoch)
            if epoch % log_frequency == 0 or epoch == num_epochs:
                save_checkpoint(experiment_directory, "latest.pth", trainer, all_lat, epoch)
                torch.save({"epoch": epoch, "loss": loss_log, "learning_rate": lr_log, "timing": timing_log,
                            "latent_magnitude": lat_mag_log, "param_magnitude": param_mag_log},
                           os.path.join(experiment_directory, logs_filename))
            logging.info("Finished {} ({}/{}) loss {:.6f} ({:.2f}s/epoch)".format(
                epoch, epoch, num_epochs, loss_log[-1] if loss_log else float("nan"), timing_log[-1]))
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()
    return {"loss": loss_log, "epochs": num_epochs}


if __name__ == "__main__":
    ap = argparse.ArgumentParser(description="Train a DeepSDF autodecoder on the fused B200 path")
    ap.add_argument("--experiment", "-e", dest="experiment_directory", required=True,
                    help="experiment directory with specs.json; checkpoints and logs are written there (reference layout)")
    ap.add_argument("--data", "-d", dest="data_source", default=None, help="overrides specs['DataSource']")
    ap.add_argument("--max-epochs", type=int, default=None)
    ap.add_argument("--strict-shuffle", action="store_true", default=None,
                    help="replicate scenes and latent codes on every GPU and shard the reference's global shuffle")
    ap.add_argument("--backend", default="nccl", help="torch.distributed backend (gloo lets several ranks share one GPU)")
    a = ap.parse_args()
    main_function(a.experiment_directory, a.data_source, a.max_epochs, a.strict_shuffle, a.backend)
