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
  OptimizerParameters/<...>.pth        {"epoch", "optimizer_state_dict": torch.optim.Adam.state_dict() of the reference's two
                                       groups (decoder, lat_vecs), all ranks' latent moments gathered}          (:106-113)
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


def gather_rows(trainer: SceneTrainer, local: torch.Tensor, num_scenes: int, world: int, rank: int) -> torch.Tensor:
    """All ranks' per-scene rows (latent codes or their Adam moments) in global scene order, on the CPU: rank r owns the
    block ``partition_scenes(num_scenes, r, world)``.  Collective in partition mode - every rank must call it."""
    if world == 1 or trainer.replicated:
        return local.detach().cpu().clone()
    import torch.distributed as dist
    most = max(len(sharding.partition_scenes(num_scenes, r, world)) for r in range(world))
    pad = torch.zeros(most, trainer.latent_size, device=trainer.device)
    pad[:local.shape[0]] = local
    parts = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(parts, pad)
    return torch.cat([p[:len(sharding.partition_scenes(num_scenes, r, world))] for r, p in enumerate(parts)], 0).cpu()


def gather_latents(trainer: SceneTrainer, num_scenes: int, world: int, rank: int) -> torch.Tensor:
    return gather_rows(trainer, trainer.lat_vecs, num_scenes, world, rank)


def adam_state_dict(trainer: SceneTrainer, latents: torch.Tensor, lat_m: torch.Tensor, lat_v: torch.Tensor, lrs) -> dict:
    """The ``state_dict()`` of the optimiser the reference builds (train_deep_sdf.py:400-411: Adam over two groups, the
    decoder's parameters in ``decoder.parameters()`` order and ``lat_vecs.weight``), filled from the trainer's flat moment
    buffers - so that ``load_optimizer`` / ``train_deep_sdf.py --continue`` (:116-131) accept the file.  Built through a real
    ``torch.optim.Adam`` so that the keys are those of the installed torch."""
    dec = trainer.decoder
    order = [p for _, p in dec.named_parameters()]                         # == DataParallel(decoder).parameters()
    flat = {id(p): (o, p.numel()) for p, o in zip(dec.flat_param_list(), trainer._offs)}
    cpu_params = [torch.nn.Parameter(p.detach().cpu().clone()) for p in order]
    lat_p = torch.nn.Parameter(latents.clone())
    opt = torch.optim.Adam([{"params": cpu_params, "lr": lrs[0]}, {"params": [lat_p], "lr": lrs[1]}])
    m, v = trainer.exp_avg.detach().cpu(), trainer.exp_avg_sq.detach().cpu()
    step = torch.tensor(float(trainer.step_count))
    for p, q in zip(order, cpu_params):
        o, n = flat[id(p)]
        opt.state[q] = {"step": step.clone(), "exp_avg": m[o:o + n].view(p.shape).clone(), "exp_avg_sq": v[o:o + n].view(p.shape).clone()}
    opt.state[lat_p] = {"step": step.clone(), "exp_avg": lat_m.clone(), "exp_avg_sq": lat_v.clone()}
    return opt.state_dict()


def load_adam_state_dict(trainer: SceneTrainer, sd: dict, lat_rows: range) -> None:
    """Inverse of ``adam_state_dict`` (also reads a file the reference's own ``save_optimizer`` wrote): per-parameter moments
    back into the flat buffers, the latent moments of THIS rank's scenes, and the step count."""
    dec = trainer.decoder
    order = [p for _, p in dec.named_parameters()]
    flat = {id(p): (o, p.numel()) for p, o in zip(dec.flat_param_list(), trainer._offs)}
    ids0, ids1 = sd["param_groups"][0]["params"], sd["param_groups"][1]["params"]
    if len(ids0) != len(order) or len(ids1) != 1:
        raise RuntimeError("optimizer state does not match this decoder (parameter groups differ)")
    st = sd["state"]
    steps = []
    for p, pid in zip(order, ids0):
        if pid not in st:
            continue
        o, n = flat[id(p)]
        trainer.exp_avg[o:o + n].copy_(st[pid]["exp_avg"].reshape(-1))
        trainer.exp_avg_sq[o:o + n].copy_(st[pid]["exp_avg_sq"].reshape(-1))
        steps.append(int(float(st[pid]["step"])))
    if ids1[0] in st:
        trainer.lat_exp_avg.copy_(st[ids1[0]]["exp_avg"][lat_rows.start:lat_rows.stop])
        trainer.lat_exp_avg_sq.copy_(st[ids1[0]]["exp_avg_sq"][lat_rows.start:lat_rows.stop])
        steps.append(int(float(st[ids1[0]]["step"])))
    trainer.set_step(max(steps) if steps else 0)


def save_checkpoint(exp: str, name: str, trainer: SceneTrainer, all_latents: torch.Tensor, all_lat_m: torch.Tensor,
                    all_lat_v: torch.Tensor, epoch: int) -> None:
    for sub in (model_params_subdir, optimizer_params_subdir, latent_codes_subdir):
        os.makedirs(os.path.join(exp, sub), exist_ok=True)
    sd = {"module." + k: v.detach().cpu().clone() for k, v in trainer.decoder.state_dict().items()}
    torch.save({"epoch": epoch, "model_state_dict": sd}, os.path.join(exp, model_params_subdir, name))
    torch.save({"epoch": epoch, "latent_codes": {"weight": all_latents}}, os.path.join(exp, latent_codes_subdir, name))
    torch.save({"epoch": epoch, "optimizer_state_dict": adam_state_dict(trainer, all_latents, all_lat_m, all_lat_v,
                                                                        trainer.learning_rates(epoch))},
               os.path.join(exp, optimizer_params_subdir, name))


def load_checkpoint(exp: str, name: str, trainer: SceneTrainer, lat_rows: range) -> int:
    """train_deep_sdf.py:426-453 (``--continue``): model, latent codes and optimiser of one saved epoch; epochs must agree."""
    fn = name + ".pth"
    model = torch.load(os.path.join(exp, model_params_subdir, fn), map_location="cpu", weights_only=False)
    lat = torch.load(os.path.join(exp, latent_codes_subdir, fn), map_location="cpu", weights_only=False)
    opt = torch.load(os.path.join(exp, optimizer_params_subdir, fn), map_location="cpu", weights_only=False)
    if not (model["epoch"] == lat["epoch"] == opt["epoch"]):
        raise RuntimeError("epoch mismatch: {} vs {} vs {}".format(model["epoch"], opt["epoch"], lat["epoch"]))
    sd = {(k[7:] if k.startswith("module.") else k): v for k, v in model["model_state_dict"].items()}
    own = trainer.decoder.state_dict()
    with torch.no_grad():                                  # the Parameters are views into the trainer's flat buffer: copy in place
        for k, v in sd.items():
            own[k].copy_(v)
    codes = lat["latent_codes"]
    w = codes["weight"] if isinstance(codes, dict) else codes.reshape(codes.shape[0], -1)        # :164-177 both layouts
    trainer.lat_vecs.copy_(w[lat_rows.start:lat_rows.stop])
    load_adam_state_dict(trainer, opt["optimizer_state_dict"], lat_rows)
    object.__setattr__(trainer.decoder, "_param_epoch", getattr(trainer.decoder, "_param_epoch", 0) + 1)
    trainer.engine.repack()
    return int(model["epoch"])


def main_function(experiment_directory: str, data_source: str = None, max_epochs: int = None,
                  strict_shuffle: bool = None, backend: str = "nccl", continue_from: str = None, batch_split: int = 1,
                  cuda_graph: bool = False) -> dict:
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
    start_epoch = 1
    if continue_from is not None:                                                                # :420-453
        logging.info('continuing from "{}"'.format(continue_from))
        saved = load_checkpoint(experiment_directory, continue_from, trainer, mine)
        logs_fn = os.path.join(experiment_directory, logs_filename)
        if os.path.isfile(logs_fn):
            lg = torch.load(logs_fn, map_location="cpu", weights_only=False)
            its_per_epoch = len(lg["loss"]) // max(1, lg["epoch"])
            loss_log, lr_log = lg["loss"][:its_per_epoch * saved], lg["learning_rate"][:saved]     # clip_logs, :203-216
            timing_log, lat_mag_log = lg["timing"][:saved], lg["latent_magnitude"][:saved]
            param_mag_log = {k: v[:saved] for k, v in lg["param_magnitude"].items()}
        start_epoch = saved + 1
    logging.info("starting from epoch {}".format(start_epoch))
    trainer.batch_split = int(batch_split)
    trainer.use_graph = bool(cuda_graph)          # partition mode, batch_split 1: every iteration is one CUDA-graph replay
    gen = torch.Generator()
    shuffle_seed = 1234 + (0 if strict_shuffle else rank)
    its = num_scenes // spb if strict_shuffle else len(dataset) // spb_local
    if world > 1 and not strict_shuffle:               # every rank must take the same number of iterations per epoch
        import torch.distributed as dist
        t = torch.tensor([its], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        its = int(t.item())
    for epoch in range(start_epoch, num_epochs + 1):
        start = time.time()
        gen.manual_seed(shuffle_seed + 7919 * epoch)       # the epoch's scene shuffle depends on the epoch only: a resumed run walks
                                                            # the batches the uninterrupted run would have walked
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
        all_lat = all_m = all_v = None
        if epoch in checkpoints or epoch % log_frequency == 0 or epoch == num_epochs:
            all_lat = gather_latents(trainer, num_scenes, world, rank)
            all_m = gather_rows(trainer, trainer.lat_exp_avg, num_scenes, world, rank)         # every rank's latent moments, not rank 0's
            all_v = gather_rows(trainer, trainer.lat_exp_avg_sq, num_scenes, world, rank)
        if rank == 0:
            lat_mag_log.append(trainer.mean_latent_magnitude())
            for n, p in trainer.decoder.named_parameters():                                      # :246-252
                param_mag_log.setdefault(n[7:] if n.startswith("module.") else n, []).append(float(p.data.norm()))
            if epoch in checkpoints:
                save_checkpoint(experiment_directory, str(epoch) + ".pth", trainer, all_lat, all_m, all_v, epoch)
            if epoch % log_frequency == 0 or epoch == num_epochs:
                save_checkpoint(experiment_directory, "latest.pth", trainer, all_lat, all_m, all_v, epoch)
                torch.save({"epoch": epoch, "loss": loss_log, "learning_rate": lr_log, "timing": timing_log,
                            "latent_magnitude": lat_mag_log, "param_magnitude": param_mag_log},
                           os.path.join(experiment_directory, logs_filename))
            logging.info("Finished {} ({}/{}) loss {:.6f} ({:.2f}s/epoch)".format(
                epoch, epoch, num_epochs, loss_log[-1] if loss_log else float("nan"), timing_log[-1]))
    trainer.release_graph()
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
    ap.add_argument("--continue", "-c", dest="continue_from", default=None,
                    help="a snapshot to continue from: 'latest' or an epoch number (train_deep_sdf.py:594-601)")
    ap.add_argument("--batch_split", dest="batch_split", type=int, default=1,
                    help="split each batch into this many chunks whose gradients are accumulated (train_deep_sdf.py:602-609): "
                         "the activation stash shrinks accordingly")
    ap.add_argument("--cuda-graph", action="store_true",
                    help="capture the optimiser step once per epoch-dependent setting and replay it (SceneTrainer.step_graphed): "
                         "same results bit for bit, no per-kernel launch gaps")
    a = ap.parse_args()
    main_function(a.experiment_directory, a.data_source, a.max_epochs, a.strict_shuffle, a.backend, a.continue_from, a.batch_split,
                  a.cuda_graph)
