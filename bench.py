#!/usr/bin/env python3
"""Headline benchmark: DeepSDF auto-decoder training step (fwd + bwd + optimiser), BASELINE.json config #2.

  python bench.py --gpus N --steps K --warmup W            this repo's sm_100a path (one process per GPU under torchrun)
  python bench.py --impl reference ...                      the CPU reference arm (oracle port of train_deep_sdf.py:483-545)

Workload: Decoder 8x512, latent 256, latent_in=[4], weight_norm, dropout 0.2 on every hidden
layer; per GPU and step 64 scenes x 16384 samples = 1 048 576 rows of synthetic (xyz, sdf) from the analytic
CornerSpheres family; clamped-L1 loss + code regulariser; Adam on decoder and on all latent rows.  A step is one full
optimiser iteration; nothing is skipped.  Weak scaling: every GPU gets its own 64 scenes per step; the flat decoder
gradient is all-reduced over NCCL.

One JSON line on stdout (rank 0).  `value` = samples/s with the batch already in HBM; `e2e` = the same step fed from
pinned host memory (H2D of the [N,4] samples inside the timed region, loss read back every step).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

FLOP_FWD_PER_ROW = 2 * 1835520            # 3.671 MFLOP/row forward (algorithmic)
FLOP_TRAIN_PER_ROW = 3 * FLOP_FWD_PER_ROW  # fwd + dgrad + wgrad
# EXECUTED tensor-core MACs per row at L = 256 (SURVEY 8(d) reconciliation rule): the latent columns of lin0 / lin4 are
# hoisted into per-scene biases, padded shapes are what the MMAs really run.
#   forward: K-chunk MMAs 512x(512*2 + 256 + 256... ) -> per layer out_pad x in_pad: lin1,2,5,6,7: 512x512; lin3: 256x512;
#            lin4: 512x256; + one K=16 aux step per layer (bias / per-scene bias / xyz): 512x16 (lin3: 256x16)
#   dgrad:   dX_l = dY_{l+1} W_{l+1}: layers 7,6,5,2,1: 512x512; lin4 (253 inputs): 256x512; lin3: 512x256
#   wgrad:   dW_l = dY_l^T X_{l-1}, l = 1..7: 512x512 each except lin4: 512x256 (lin3's 253 outputs ride in a 512-wide block)
EXEC_MAC_FWD = 5 * 512 * 512 + 256 * 512 + 512 * 256 + 7 * 512 * 16 + 256 * 16
EXEC_MAC_DGRAD = 5 * 512 * 512 + 256 * 512 + 512 * 256
EXEC_MAC_WGRAD = 6 * 512 * 512 + 512 * 256
EXEC_FLOP = {"fused_forward": 2 * EXEC_MAC_FWD, "fused_dgrad": 2 * EXEC_MAC_DGRAD, "wgrad": 2 * EXEC_MAC_WGRAD}
SCENES_PER_STEP = 64
SAMPLES_PER_SCENE = 16384
SCENES_TOTAL = 1024                        # lat_vecs rows per GPU
N_HOST_BATCHES = 4

SPECS = {
    "Description": "bench: original DeepSDF decoder 8x512, latent 256 (BASELINE.json config #2)",
    "NetworkArch": "deep_sdf_decoder",
    "NetworkSpecs": {
        "dims": [512] * 8, "dropout": list(range(8)), "dropout_prob": 0.2, "norm_layers": list(range(8)),
        "latent_in": [4], "xyz_in_all": False, "use_tanh": False, "latent_dropout": False, "weight_norm": True,
        "geom_dimension": 3,
    },
    "CodeLength": 256,
    "LearningRateSchedule": [
        {"Type": "Step", "Initial": 0.0005, "Interval": 500, "Factor": 0.5},
        {"Type": "Step", "Initial": 0.001, "Interval": 500, "Factor": 0.5},
    ],
    "SamplesPerScene": SAMPLES_PER_SCENE, "ScenesPerBatch": SCENES_PER_STEP, "ClampingDistance": 0.1,
    "CodeRegularization": True, "CodeRegularizationLambda": 1e-4, "CodeBound": 1.0,
}


# dram__bytes_read.sum + dram__bytes_write.sum per launch of each tensor kernel at this workload: a STATIC copy of the
# committed ncu capture (profiles/r02_ncu_v3_summary.txt), not re-measured by this script; algorithmic stash traffic:
# forward 8.6 GB, dgrad 7.5 GB, wgrad 14.7 GB
NCU_DRAM_BYTES = {"fused_forward": 8.562e9, "fused_dgrad": 7.708e9, "wgrad": 14.510e9}


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return {"burst": d.get("bf16_tflops"), "sustained": d.get("bf16_tflops_sustained"), "hbm": d.get("hbm_gbs"),
                "source": "measured"}
    return {"burst": 1590.0, "sustained": 1400.0, "hbm": 6650.0, "source": "fallback"}


# ------------------------------------------------------------------------------------------------ synthetic data
def corner_spheres_sdf_torch(xyz: torch.Tensor, radius: torch.Tensor, limit: float = 0.9) -> torch.Tensor:
    """sdf_sampler/microstructures.py:23-36 CornerSpheresSDF (inf-norm box minus eight 3-norm corner 'spheres')."""
    out = xyz.abs().amax(dim=1) - limit
    for sx in (-1.0, 1.0):
        for sy in (-1.0, 1.0):
            for sz in (-1.0, 1.0):
                c = torch.tensor([sx, sy, sz], device=xyz.device)
                d = (xyz - c).abs().pow(3).sum(dim=1).pow(1.0 / 3.0) - radius
                out = torch.maximum(out, -d)
    return out


def cross_ms_sdf_torch(xyz: torch.Tensor, radius: torch.Tensor) -> torch.Tensor:
    """sdf_sampler/microstructures.py:3-20 CrossMsSDF (inf-norm 'box' united with three axis cylinders)."""
    out = xyz.abs().amax(dim=-1)
    for a, b in ((1, 2), (0, 2), (0, 1)):
        out = torch.minimum(out, torch.sqrt(xyz[..., a] ** 2 + xyz[..., b] ** 2) - radius)
    return out


def make_batch(gen: torch.Generator, device, scene_ids: torch.Tensor, scenes_total: int) -> torch.Tensor:
    """[S*SPS, 4] fp32 (x, y, z, sdf), scene-major, as the reference's DataLoader yields after reshape(-1, 4)."""
    S = scene_ids.numel()
    n = S * SAMPLES_PER_SCENE
    xyz = torch.rand(n, 3, generator=gen, device=device) * 2.0 - 1.0
    r = (0.4 + 0.6 * scene_ids.to(device).float() / max(1, scenes_total - 1)).repeat_interleave(SAMPLES_PER_SCENE)
    sdf = corner_spheres_sdf_torch(xyz, r)
    return torch.cat([xyz, sdf.unsqueeze(1)], dim=1).contiguous()


# ------------------------------------------------------------------------------------------------ clocks
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self.proc, self.thread = index, [], None, None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None
            return
        self.thread = threading.Thread(target=self._read, daemon=True)
        self.thread.start()

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:  # noqa: BLE001
            self.proc.kill()
        sm, mx, reasons, pw = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for row in self.rows:
            f = [x.strip() for x in row.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1])); pw.append(float(f[2]))
            except ValueError:
                continue
            for nm, val in zip(names, f[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(nm)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        hot = [c for c, p in zip(sm, pw) if p >= 0.5 * max(pw)] or sm
        return {"sm_mhz": statistics.median(hot), "sm_max_mhz": max(mx), "power_w_max": max(pw), "samples": len(sm),
                "reasons": sorted(reasons)}


# ------------------------------------------------------------------------------------------------ CPU reference arm
def cpu_train_step_rate(n_scenes: int, steps: int, warmup: int):
    """Oracle port of the reference train step (train_deep_sdf.py:483-545 with the reference Decoder's math) on the host
    cores: same network, same per-row work, ``n_scenes`` x 16384 rows per step.  Returns (samples/s, ms/step, threads)."""
    from oracle import decoder_oracle as O
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    ns = SPECS["NetworkSpecs"]
    spec = O.OracleSpec(latent_size=SPECS["CodeLength"], **ns)
    torch.manual_seed(0)
    params = O.make_params(spec, 0)
    n = n_scenes * SAMPLES_PER_SCENE
    g = torch.Generator().manual_seed(2)
    table = torch.randn(SCENES_TOTAL, spec.latent_size, generator=g) / (spec.latent_size ** 0.5)
    ids = torch.arange(n_scenes)
    batch = make_batch(g, "cpu", ids, SCENES_TOTAL)
    row_scene = ids.repeat_interleave(SAMPLES_PER_SCENE)
    names = list(params)
    m = {k: torch.zeros_like(v) for k, v in params.items()}
    v2 = {k: torch.zeros_like(v) for k, v in params.items()}
    lm, lv = torch.zeros_like(table), torch.zeros_like(table)
    times = []
    for it in range(warmup + steps):
        t0 = time.perf_counter()
        masks = {l: (torch.rand(n, spec.out_dims[l], generator=g) >= 0.2).float() for l in range(spec.num_linear - 1)}
        _, grads, lat_grad, _ = O.train_loss_and_grads(spec, params, table, row_scene, batch[:, :3], batch[:, 3], 0.1, 1e-4, 1,
                                                       masks=masks, code_bound=1.0)
        for k in names:
            O.adam_step_(params[k], grads[k], m[k], v2[k], it + 1, 5e-4)
        O.adam_step_(table, lat_grad, lm, lv, it + 1, 1e-3)
        dt = time.perf_counter() - t0
        if it >= warmup:
            times.append(dt)
    ms = 1e3 * sum(times) / len(times)
    return n / (ms / 1e3), ms, threads


def cpu_grid_rate(n_grid: int):
    """Oracle port of create_mesh's chunk loop (deep_sdf/mesh.py:37-73, 32768-row chunks, no marching cubes) on the host
    cores at a bounded resolution: the CPU figure printed beside the 512^3 grid queries/s."""
    from oracle import decoder_oracle as O
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    spec = O.OracleSpec(latent_size=SPECS["CodeLength"], **SPECS["NetworkSpecs"])
    params = O.make_params(spec, 0)
    g = torch.Generator().manual_seed(1)
    lat = torch.randn(spec.latent_size, generator=g) / (spec.latent_size ** 0.5)
    O.create_mesh_sdf(spec, params, lat, 16)                  # warm-up
    t0 = time.perf_counter()
    O.create_mesh_sdf(spec, params, lat, n_grid)
    dt = time.perf_counter() - t0
    return {"queries_per_s": n_grid ** 3 / dt, "cores": threads, "kind": "port", "ms": 1e3 * dt,
            "sample": f"N = {n_grid} grid ({n_grid ** 3} queries) through the oracle's create_mesh loop, SDF only, PyTorch CPU fp32"}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    n_scenes = 4
    rate, ms, threads = cpu_train_step_rate(n_scenes, args.steps, args.warmup)
    sample = f"{n_scenes} scenes x {SAMPLES_PER_SCENE} rows per step ({n_scenes * SAMPLES_PER_SCENE} of the 1048576-row batch), same network and step math"
    line = {
        "impl": "reference", "metric": "train SDF samples/s (fwd+bwd+optimizer)", "value": rate, "unit": "samples/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "config#2: Decoder 8x512 L=256 latent_in=[4], 64 scenes x 16384 samples/step/GPU (CPU arm: bounded sample)",
                   "note": "reference arm = oracle port of train_deep_sdf.py:483-545 (PyTorch CPU fp32); the reference is Python and does not travel to the GPU box"},
        "cpu_baseline": {"value": rate, "unit": "samples/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": rate, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------ GPU arm
def _teardown(dist) -> None:
    """destroy_process_group with a watchdog: the measurement is complete and printed; a communicator teardown that does not
    return within a minute must not hold the launcher (exit code 0 either way)."""
    import threading
    t = threading.Timer(60.0, lambda: os._exit(0))
    t.daemon = True
    t.start()
    torch.cuda.synchronize()
    dist.barrier()
    dist.destroy_process_group()
    t.cancel()


def run_gpu(args):
    import torch.distributed as dist
    from deepsdf_b200.trainer import SceneTrainer

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (sm_100a); there is no CPU path for the product arm")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    peaks = load_peaks()

    from deepsdf_b200 import _lib as L
    lib = L.lib()
    strong = args.scaling == "strong"
    if SCENES_PER_STEP % world:
        raise SystemExit("strong scaling splits the 64-scene batch: --gpus must divide 64")
    S_step = SCENES_PER_STEP // world if strong else SCENES_PER_STEP      # scenes this rank steps per iteration
    trainer = SceneTrainer(SPECS, SCENES_TOTAL, dev, rank=rank, world_size=world, seed=0)
    N = S_step * SAMPLES_PER_SCENE
    gen = torch.Generator(device=dev).manual_seed(2 + rank)
    gcpu = torch.Generator().manual_seed(5 + rank)
    dev_batches, host_batches, dev_ids, host_ids = [], [], [], []
    for b in range(N_HOST_BATCHES):
        ids = torch.randperm(SCENES_TOTAL, generator=gcpu)[:S_step].sort().values
        db = make_batch(gen, dev, ids, SCENES_TOTAL)
        dev_batches.append(db)
        host_batches.append(db.cpu().pin_memory())
        dev_ids.append(ids.to(dev))
        host_ids.append(ids.pin_memory())
    global_rows = N * world
    row0 = rank * N

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms: float) -> float:
        if world == 1:
            return ms
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- warm-up
    for i in range(max(3, args.warmup)):
        trainer.step(dev_batches[i % N_HOST_BATCHES], dev_ids[i % N_HOST_BATCHES], epoch=1, global_rows=global_rows, row0=row0)
    barrier()

    # ---- region 0 (eager launches, a few steps): per-kernel CUDA events on the launching stream + the eager ms/step
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    n_eager = min(args.steps, 10)
    trainer.engine.kernel_events = []
    a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    launches0 = int(lib.dsdf_launch_count())
    a0.record()
    for i in range(n_eager):
        loss = trainer.step(dev_batches[i % N_HOST_BATCHES], dev_ids[i % N_HOST_BATCHES], epoch=1, global_rows=global_rows,
                            row0=row0)
    a1.record()
    barrier()
    launches_eager = (int(lib.dsdf_launch_count()) - launches0) // n_eager
    ms_eager = max_over_ranks(a0.elapsed_time(a1)) / n_eager
    kev = trainer.engine.kernel_events
    trainer.engine.kernel_events = None

    # ---- timed region 1: inputs resident in HBM; the step is ONE CUDA-graph replay (SceneTrainer.step_graphed: the same
    #      kernels, bit-identical results - tests/test_train_features_gpu.py) unless --no-graph
    graphed = not args.no_graph
    step_fn = trainer.step_graphed if graphed else trainer.step
    for i in range(2):
        step_fn(dev_batches[i % N_HOST_BATCHES], dev_ids[i % N_HOST_BATCHES], epoch=1, global_rows=global_rows, row0=row0)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    launches0 = int(lib.dsdf_launch_count())
    e0.record()
    for i in range(args.steps):
        loss = step_fn(dev_batches[i % N_HOST_BATCHES], dev_ids[i % N_HOST_BATCHES], epoch=1, global_rows=global_rows, row0=row0)
    e1.record()
    barrier()
    # this repo's kernels inside timed region 1: counted by the library at its launch statements (eager) / at capture (graph)
    gpu_launches = trainer.graph_launches * args.steps if graphed else int(lib.dsdf_launch_count()) - launches0
    ms_total = max_over_ranks(e0.elapsed_time(e1))
    ms_step = ms_total / args.steps
    value = global_rows / (ms_step / 1e3)
    loss_dev = [float(x) for x in loss.cpu()]

    # ---- timed region 2: end to end from pinned host memory, loss read back every step.  The user-facing loop of a host-fed
    #      run: HostFeeder.submit (H2D copy of the NEXT batch on a copy stream, what DataLoader workers + .cuda() do in the
    #      reference) then HostFeeder.step on the current one; every timed step contains one 16.8 MB copy and one loss read (asynchronous, two steps behind: HostFeeder.step_async).
    from deepsdf_b200.trainer import HostFeeder
    LAG = 2                               # steps the host runs ahead of the loss it reads (absorbs host-side jitter between ranks)
    feeder = HostFeeder(trainer, N, S_step, depth=LAG + 2, graphed=graphed)
    feeder.submit(host_batches[0], host_ids[0])
    for i in range(2):
        feeder.submit(host_batches[(i + 1) % N_HOST_BATCHES], host_ids[(i + 1) % N_HOST_BATCHES])
        feeder.step_async(1, global_rows=global_rows, row0=row0).wait()
    barrier()
    e2, e3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e2.record()
    last, pending = None, []
    for i in range(args.steps):
        nb = (i + 3) % N_HOST_BATCHES
        feeder.submit(host_batches[nb], host_ids[nb])          # next batch's copy runs under this step
        pending.append(feeder.step_async(1, global_rows=global_rows, row0=row0))
        if len(pending) > LAG:
            last = pending.pop(0).wait()  # device -> host read of an EARLIER step's loss, while later steps are already queued
    while pending:
        last = pending.pop(0).wait()      # every step's loss has been read back inside the timed region
    e3.record()
    barrier()
    ms_e2e = max_over_ranks(e2.elapsed_time(e3)) / args.steps
    clk = clocks.stop() if rank == 0 else None
    e2e_value = global_rows / (ms_e2e / 1e3)
    h2d = host_batches[0].numel() * 4 + host_ids[0].numel() * 8
    d2h = 8

    # ---- per-kernel durations (CUDA events on the launching stream, inside timed region 1)
    per = {}
    for name, a, b in kev:
        per.setdefault(name, []).append(a.elapsed_time(b))
    kern = {k: sum(v) / len(v) for k, v in per.items()}
    tensor_kernels = {k: kern[k] for k in ("fused_forward", "fused_dgrad", "wgrad") if k in kern}
    dom = max(tensor_kernels, key=tensor_kernels.get)
    dom_ms = tensor_kernels[dom]
    achieved = FLOP_FWD_PER_ROW * N / (dom_ms / 1e3) / 1e12
    peak = peaks["sustained"]
    roofline = {
        "bound": "tensor", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak,
        "traffic": NCU_DRAM_BYTES.get(dom), "traffic_unit": "bytes per launch (dram read + write, ncu --set full, profiles/r02_ncu_v3_summary.txt)",
        "peak_source": f"{peaks['source']} bf16 sustained (kernel timed inside a long step)",
        "flops_per_launch": FLOP_FWD_PER_ROW * N,
        "kernels_ms": kern,
        "kernels_tflops": {k: FLOP_FWD_PER_ROW * N / (v / 1e3) / 1e12 for k, v in tensor_kernels.items()},
        # SURVEY 8(d) reconciliation: the tensor-core work the kernels actually issue (latent columns hoisted into per-scene
        # biases, padded shapes) - what ncu's tensor-pipe % reflects; never to be read as the algorithmic figure above
        "executed_tflops": EXEC_FLOP[dom] * N / (dom_ms / 1e3) / 1e12,
        "executed_frac": EXEC_FLOP[dom] * N / (dom_ms / 1e3) / 1e12 / peak,
        "kernels_executed_tflops": {k: EXEC_FLOP[k] * N / (v / 1e3) / 1e12 for k, v in tensor_kernels.items()},
        "step_executed_frac_of_peak": sum(EXEC_FLOP.values()) * N / (ms_step / 1e3) / 1e12 / peak,
        "traffic_source": "static (ncu capture of the same launch, committed under profiles/)",
        "step_tflops_algorithmic": FLOP_TRAIN_PER_ROW * N / (ms_step / 1e3) / 1e12,
        "step_frac_of_peak": FLOP_TRAIN_PER_ROW * N / (ms_step / 1e3) / 1e12 / peak,
        "step_frac_of_datasheet_2250": FLOP_TRAIN_PER_ROW * N / (ms_step / 1e3) / 1e12 / 2250.0,
    }

    # ---- strong scaling ("each rank's batch = S/G scenes"): the FIXED 64 x 16384 batch split over the ranks,
    #      global-N loss normaliser, one all-reduce of the decoder gradient per step.  Reported as an extra object of the
    #      default (weak) line so that the driver's 1/2/4/8-GPU SCALE record carries it.  On one GPU the per-rank shards of
    #      2 / 4 / 8 GPUs are timed alone (no collective): the compute-side bound T1 / (n * T_shard) of the strong efficiency.
    def timed(scenes, steps, g_rows, r0):
        n_rows = scenes * SAMPLES_PER_SCENE
        sl = [(b[:n_rows], i[:scenes].contiguous()) for b, i in zip(dev_batches, dev_ids)]
        for k in range(3):
            step_fn(sl[k % N_HOST_BATCHES][0], sl[k % N_HOST_BATCHES][1], epoch=1, global_rows=g_rows, row0=r0)
        barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for k in range(steps):
            step_fn(sl[k % N_HOST_BATCHES][0], sl[k % N_HOST_BATCHES][1], epoch=1, global_rows=g_rows, row0=r0)
        b.record()
        barrier()
        return max_over_ranks(a.elapsed_time(b)) / steps

    strong_obj = None
    if not strong:
        g_rows = SCENES_PER_STEP * SAMPLES_PER_SCENE
        if world > 1 and SCENES_PER_STEP % world == 0:
            s_loc = SCENES_PER_STEP // world
            ms_s = timed(s_loc, args.steps, g_rows, rank * s_loc * SAMPLES_PER_SCENE)
            strong_obj = {"value": g_rows / (ms_s / 1e3), "unit": "samples/s", "ms_per_step": ms_s, "scenes_per_gpu": s_loc,
                          "global_rows_per_step": g_rows, "n_gpus": world,
                          "what": "fixed 64 x 16384 batch split over the GPUs, decoder-gradient all-reduce every step"}
        elif world == 1:
            shards = {}
            for n_g in (2, 4, 8):
                shards[str(n_g)] = timed(SCENES_PER_STEP // n_g, args.steps, g_rows, 0)
            strong_obj = {"value": value, "unit": "samples/s", "ms_per_step": ms_step, "scenes_per_gpu": SCENES_PER_STEP, "n_gpus": 1,
                          "global_rows_per_step": g_rows, "shard_ms_per_step": shards,
                          "shard_efficiency_bound": {k: ms_step / (int(k) * v) for k, v in shards.items()},
                          "what": "one GPU: the whole batch (= the headline) and, alone on this GPU, the per-rank shards of 2/4/8 GPUs "
                                  "(compute only, no collective): T1 / (n * T_shard) bounds the strong-scaling efficiency from above"}

    # ---- secondary metrics of BASELINE.json (never part of the timed train region): grid queries/s (config #4) and the
    #      batched latent-only optimisation (config #3), each sharded without a collective
    extras = {}
    if not args.no_extras:
        try:
            extras = run_extras(trainer, dev, rank, world, barrier, max_over_ranks)
        except Exception as exc:  # noqa: BLE001 - the headline line must still be printed
            extras = {"error": repr(exc)[:300]}

    trainer.release_graph()           # a live graph holding NCCL kernels blocks the communicator's teardown
    if rank != 0:
        if world > 1:
            _teardown(dist)
        return

    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        n_sc = 2
        rate, ms_cpu, threads = cpu_train_step_rate(n_sc, steps=2, warmup=1)
        cpu = {"value": rate, "unit": "samples/s", "cores": threads, "kind": "port",
               "sample": f"{n_sc} scenes x {SAMPLES_PER_SCENE} rows per step, 1 warm-up + 2 timed steps, oracle port of train_deep_sdf.py:483-545 (PyTorch CPU fp32)",
               "ms_per_step": ms_cpu}
        if isinstance(extras.get("grid"), dict):
            extras["grid"]["cpu_baseline"] = cpu_grid_rate(64)

    line = {
        "metric": "train SDF samples/s (fwd+bwd+optimizer)", "value": value, "unit": "samples/s", "n_gpus": world,
        "steps": args.steps, "warmup": max(3, args.warmup), "ms_per_step": ms_step, "higher_is_better": True,
        "scaling": "strong" if strong else "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": "config#2: Decoder 8x512 L=256 latent_in=[4] weight_norm dropout0.2, " +
                               (f"64 scenes x 16384 samples per step split over {world} GPU(s)" if strong else
                                "64 scenes x 16384 samples per GPU per step"),
                   "rows_per_gpu_step": N, "global_rows_per_step": global_rows, "parallelism": f"dp{world}",
                   "exchange": ("none (one GPU)" if world == 1 else
                                "own NVLS all-reduce kernel (multimem.ld_reduce/st over NVSwitch multicast, dsdf_nvls_allreduce_f32)"
                                if trainer._nvls is not None else "ncclAllReduce"),
                   "cache": "per-step working set (bf16 activation + dY stashes, 17 GB) is far larger than the 126 MB L2; 4 rotating input batches"},
        "e2e": {"value": e2e_value, "unit": "samples/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": ms_e2e},
        "gpu_launches": gpu_launches,         # this repo's kernels inside timed region 1 (dsdf_launch_count; NCCL's are not ours)
        "launch_mode": {"timed_region": "cuda_graph" if graphed else "eager", "kernels_per_step": gpu_launches // args.steps,
                        "eager_ms_per_step": ms_eager, "eager_kernels_per_step": launches_eager,
                        "what": "eager = the same step launched kernel by kernel (region 0, also the source of kernels_ms)"},
        "roofline": roofline,
        "cpu_baseline": cpu,
        "clocks": clk,
        "loss_last_step": loss_dev,
        "loss_last_step_e2e": [float(x) for x in last],
        # secondary metrics LAST (driver tails keep the end of the line): sampler, iso-surface, microstructure consumers,
        # config #5, config #3 (latent_opt), config #4 (grid), strong scaling
        "sampler": extras.get("sampler"), "isosurface": extras.get("isosurface"), "microstructure": extras.get("microstructure"),
        "small_nets": extras.get("small_nets"), "compat": extras.get("compat"), "extras_error": extras.get("error"),
        "config5": extras.get("config5"), "latent_opt": extras.get("latent_opt"), "grid": extras.get("grid"),
        "strong": strong_obj,
    }
    print(json.dumps(line), flush=True)
    if world > 1:
        _teardown(dist)


def run_extras(trainer, dev, rank, world, barrier, max_over_ranks):
    """BASELINE.json configs #4 and #3 on the trained-shape decoder of the bench (random weights):
    grid   create_mesh's 512^3 grid (mesh.py:37-73), z-slab sharded over the ranks, SDF only and SDF + full input gradient
           (dSDF/dxyz [n,3] and dSDF/dlatent [n,256] written to HBM chunk by chunk);
    latent reconstruct.py-style latent-only optimisation, 1024 shapes per GPU x 8000 samples, a bounded number of iterations."""
    from deepsdf_b200.mesh import slab_for_rank
    from deepsdf_b200.reconstruct import LatentOptimizer
    from deepsdf_b200.utils import fused_engine_for
    out = {}
    dec = trainer.decoder
    dec.eval()
    trainer.engine._bufs, trainer.engine._bufs_key = {}, None     # release the 18 GB training stash
    torch.cuda.empty_cache()
    eng = fused_engine_for(dec)
    Lz = eng.spec.latent_size
    NG = 512
    i0, i1 = slab_for_rank(NG, rank, world)
    first, total = i0 * NG * NG, (i1 - i0) * NG * NG
    chunk = 1 << 21
    xyz = torch.zeros(chunk, 3, device=dev)
    ts = torch.zeros(chunk // 128, dtype=torch.int32, device=dev)
    sdf = torch.empty(chunk, device=dev)
    G = torch.empty(chunk, Lz + 3, device=dev)
    cb = eng.scene_bias(trainer.lat_vecs[:1].contiguous(), torch.zeros(1, dtype=torch.int64, device=dev))

    cbx = eng.scene_bias_exact(trainer.lat_vecs[:1].contiguous(), torch.zeros(1, dtype=torch.int64, device=dev))

    def sweep(mode):
        head = 0
        while head < total:
            n = min(chunk, total - head)
            eng.grid_points(NG, first + head, n, out=xyz)
            if mode == "sdf_and_grad":
                eng.sdf_and_input_grad(cb, ts, xyz, G=G, sdf=sdf)
            elif mode == "sdf_exact":
                eng.decode_scenes(cbx, ts, xyz, out=sdf, exact=True)      # split-fp16 mode: <= 1e-5 vs fp32 (csrc/chain_exact.cu)
            else:
                eng.decode_scenes(cb, ts, xyz, out=sdf)
            head += n

    res = {}
    for mode in ("sdf_only", "sdf_and_grad", "sdf_exact"):
        sweep(mode)                                          # warm-up (allocations, attribute setup)
        barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); sweep(mode); b.record()
        barrier()
        ms = max_over_ranks(a.elapsed_time(b))
        res[mode] = {"ms": ms, "queries_per_s": NG ** 3 / (ms / 1e3)}
    res["sdf_exact"]["what"] = "high-precision tensor-core mode: (hi, lo) fp16 operand pairs, <= 1e-5 vs the fp32 arithmetic (tests/test_exact_gpu.py)"
    flop = 2 * FLOP_FWD_PER_ROW
    out["grid"] = {"N": NG, "queries": NG ** 3, "sharding": f"z-slab x{world}", "chunk_rows": chunk, **res,
                   "sdf_and_grad_tflops_algorithmic": flop * NG ** 3 / (res["sdf_and_grad"]["ms"] / 1e3) / 1e12,
                   "outputs": f"sdf [n], G [n,{Lz + 3}] = dSDF/d[latent, xyz] fp32 per chunk"}
    del xyz, ts, sdf, G
    eng._bufs, eng._bufs_key = {}, None
    torch.cuda.empty_cache()

    # iso-surface of a device-resident 512^3 volume (the tail of create_mesh, mesh.py:88-155): analytic sphere, rank 0's GPU
    try:
        from deepsdf_b200.mesh import extract_isosurface
        gl = torch.linspace(-1, 1, NG, device=dev)
        vol = (torch.sqrt(gl[:, None, None] ** 2 + gl[None, :, None] ** 2 + gl[None, None, :] ** 2) - 0.6).contiguous()
        tri = extract_isosurface(vol, (-1, -1, -1), 2.0 / (NG - 1))
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); tri = extract_isosurface(vol, (-1, -1, -1), 2.0 / (NG - 1)); b.record()
        torch.cuda.synchronize()
        ms = a.elapsed_time(b)
        out["isosurface"] = {"N": NG, "ms": ms, "triangles": int(tri.shape[0]), "voxels_per_s": NG ** 3 / (ms / 1e3),
                             "volume_GBps": 2 * vol.numel() * 4 / (ms / 1e3) / 1e9,
                             "what": "marching tetrahedra, count + scan + emit, volume read twice (algorithmic bytes = 2 x 4 B/voxel)"}
        del vol, tri, gl
        torch.cuda.empty_cache()
    except Exception as exc:  # noqa: BLE001
        out["isosurface"] = {"error": repr(exc)[:200]}

    # microstructure evaluation with per-vertex codes + d sdf / d latent (mesh.py:457-514, 405-422): L = 2 network of config #5
    try:
        from deepsdf_b200.mesh import evaluate_network_sdf
        from deepsdf_b200.networks.deep_sdf_decoder import Decoder as _Dec
        torch.manual_seed(5)
        ns2 = dict(SPECS["NetworkSpecs"])
        dec2 = _Dec(2, **ns2).to(dev).eval()
        n_v = 450000
        so = (torch.rand(n_v, 3, device=dev) - 0.5) * 2.1
        lv = torch.randn(n_v, 2, device=dev) * 0.3
        capd = {"x0": {"cap": 1, "measure": 0.1}, "x1": {"cap": 1, "measure": 0.1}}
        evaluate_network_sdf(dec2, lv, so, [2, 2, 2], capd, with_grad=True)
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(5):
            evaluate_network_sdf(dec2, lv, so, [2, 2, 2], capd, with_grad=True)
        b.record()
        torch.cuda.synchronize()
        ms = a.elapsed_time(b) / 5
        out["microstructure"] = {"vertices": n_v, "latent_size": 2, "ms": ms, "vertices_per_s": n_v / (ms / 1e3),
                                 "what": "tiling transform + SDF with per-vertex codes + caps + d sdf/d latent for ALL control points "
                                         "(one G instead of the L x n_cpt jvp loop: 0.8-5.8 s per call in the reference's logs)"}
        del dec2, so, lv
        torch.cuda.empty_cache()
    except Exception as exc:  # noqa: BLE001
        out["microstructure"] = {"error": repr(exc)[:200]}

    # the networks the reference's optimisation runs actually use (optimization_runs/*/config.json:8), at the point counts its logs
    # record: evaluate_network_sdf = tiling transform + SDF + caps + d sdf / d latent for every vertex,
    # ONE fused fp32 launch of csrc/smallnet.cu instead of the per-(latent, control point) jvp loop
    try:
        from deepsdf_b200.mesh import evaluate_network_sdf
        from deepsdf_b200.networks.deep_sdf_decoder import Decoder as _Dec
        small = {}
        for tag, Lz, dims, lat_in, n_v, ref_s in (("4x32_L2", 2, [32] * 4, [2], 316030, "0.90-1.17 s"),
                                                  ("6x128_L1", 1, [128] * 6, [2], 450241, "0.80-5.8 s")):
            torch.manual_seed(7)
            d3 = _Dec(Lz, dims=dims, geom_dimension=3, dropout=list(range(8)), dropout_prob=0.2, norm_layers=list(range(8)),
                      latent_in=lat_in, weight_norm=True, xyz_in_all=False, use_tanh=False, latent_dropout=False).to(dev).eval()
            so = (torch.rand(n_v, 3, device=dev) - 0.5) * 2.1
            lv = torch.randn(n_v, Lz, device=dev) * 0.3
            capd = {"x0": {"cap": 1, "measure": 0.1}, "x1": {"cap": 1, "measure": 0.1}}
            evaluate_network_sdf(d3, lv, so, [2, 2, 2], capd, with_grad=True)
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(10):
                evaluate_network_sdf(d3, lv, so, [2, 2, 2], capd, with_grad=True)
            b.record()
            torch.cuda.synchronize()
            ms = a.elapsed_time(b) / 10
            small[tag] = {"vertices": n_v, "ms": ms, "vertices_per_s": n_v / (ms / 1e3), "reference_log_per_call": ref_s + " (A100, jvp loop)"}
            del d3, so, lv
        out["small_nets"] = small
        torch.cuda.empty_cache()
    except Exception as exc:  # noqa: BLE001
        out["small_nets"] = {"error": repr(exc)[:200]}

    # the drop-in module's own boundary call, decoder(input[N, L+3]) (what unchanged reference scripts execute): exact fp32
    # CUDA-core mode at config #2's network (fwd + bwd with parameter gradients), and the eval-mode tensor-core route (L = 16)
    try:
        from deepsdf_b200.networks.deep_sdf_decoder import Decoder as _Dec2
        comp = {}
        n_c = 1 << 16
        torch.manual_seed(9)
        d4 = _Dec2(256, **SPECS["NetworkSpecs"]).to(dev).train()
        x4 = torch.cat([torch.randn(n_c, 256, device=dev) / 16, torch.rand(n_c, 3, device=dev) * 2 - 1], 1)
        for _ in range(2):
            d4.zero_grad(); d4(x4).sum().backward()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(3):
            d4.zero_grad(); d4(x4).sum().backward()
        b.record()
        torch.cuda.synchronize()
        ms = a.elapsed_time(b) / 3
        comp["fp32_generic_train_L256"] = {"rows": n_c, "ms_fwd_bwd": ms, "samples_per_s": n_c / (ms / 1e3),
                                           "what": "Decoder.forward + backward (parameter gradients), fp32 CUDA-core kernels: the exact (1e-5) mode"}
        del x4
        d4.eval()
        n_b = 1 << 19
        x6 = torch.cat([torch.randn(n_b, 256, device=dev) / 16, torch.rand(n_b, 3, device=dev) * 2 - 1], 1)
        with torch.no_grad():
            d4(x6)
            torch.cuda.synchronize()
            a.record()
            for _ in range(5):
                d4(x6)
            b.record()
            torch.cuda.synchronize()
        ms_b = a.elapsed_time(b) / 5
        xg = x6.clone().requires_grad_(True)
        (g6,) = torch.autograd.grad(d4(xg).sum(), xg)
        torch.cuda.synchronize()
        a.record()
        (g6,) = torch.autograd.grad(d4(xg).sum(), xg)
        b.record()
        torch.cuda.synchronize()
        comp["eval_L256"] = {"rows": n_b, "ms_forward": ms_b, "queries_per_s": n_b / (ms_b / 1e3),
                             "ms_forward_and_input_grad": a.elapsed_time(b),
                             "what": "decoder(input[N, 259]) in eval mode: per-row codes as bf16 K chunks of the tcgen05 chain; "
                                     "input.grad from the fused dgrad chain + input-gradient GEMM"}
        del d4, x6, xg, g6
        d5 = _Dec2(16, **SPECS["NetworkSpecs"]).to(dev).eval()
        n_e = 1 << 20
        x5 = torch.cat([torch.randn(n_e, 16, device=dev) / 4, torch.rand(n_e, 3, device=dev) * 2 - 1], 1)
        with torch.no_grad():
            d5(x5)
            torch.cuda.synchronize()
            a.record()
            for _ in range(5):
                d5(x5)
            b.record()
            torch.cuda.synchronize()
            ms_f = a.elapsed_time(b) / 5
            d5.fused_eval = False
            d5(x5[:n_c])
            torch.cuda.synchronize()
            a.record()
            d5(x5[:n_c])
            b.record()
            torch.cuda.synchronize()
            ms_g = a.elapsed_time(b)
        comp["eval_L16"] = {"rows": n_e, "ms_tensor_core_rows_mode": ms_f, "queries_per_s": n_e / (ms_f / 1e3),
                            "fp32_generic_queries_per_s": n_c / (ms_g / 1e3),
                            "what": "decoder(input[N, 19]) in eval mode: bf16 tcgen05 chain (row-latent mode) vs the fp32 generic path"}
        out["compat"] = comp
        del d5, x5
        torch.cuda.empty_cache()
    except Exception as exc:  # noqa: BLE001
        out["compat"] = {"error": repr(exc)[:200]}

    # device-resident dataset + sampler (deep_sdf/data.py:75-110 on the GPU): 64 scenes x 16384 of 100000 stored rows
    from deepsdf_b200.data import DeviceSDFSamples
    n_st = 50000
    arrays = []
    g2 = torch.Generator(device=dev).manual_seed(3 + rank)
    for s in range(SCENES_PER_STEP):
        blk = make_batch(g2, dev, torch.tensor([s]), SCENES_TOTAL)
        extra = make_batch(g2, dev, torch.tensor([s]), SCENES_TOTAL)
        rows = torch.cat([blk, extra, make_batch(g2, dev, torch.tensor([s]), SCENES_TOTAL)], 0)[:2 * n_st].cpu().numpy()
        pos, neg = rows[rows[:, 3] > 0], rows[rows[:, 3] <= 0]
        arrays.append((pos, neg))
    try:
        ds = DeviceSDFSamples(None, None, SAMPLES_PER_SCENE, 3, dev, arrays=arrays)
        ids_s = torch.arange(SCENES_PER_STEP, device=dev)
        buf = ds.sample_batch(ids_s, 1)
        barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for i in range(20):
            ds.sample_batch(ids_s, 2 + i, out=buf)
        b.record()
        barrier()
        ms = max_over_ranks(a.elapsed_time(b)) / 20
        out["sampler"] = {"ms_per_batch": ms, "rows_per_s": world * buf.shape[0] / (ms / 1e3), "stored_rows_per_gpu": ds.total_rows,
                          "what": "on-device half-positive/half-negative subsample without replacement of a 64-scene batch"}
        del ds, buf
    except Exception as exc:  # noqa: BLE001 - e.g. a scene without enough rows of one sign in the synthetic family
        out["sampler"] = {"error": repr(exc)[:200]}
    S, ns, M, iters = 1024, 8000, 16384, 800
    gen = torch.Generator(device=dev).manual_seed(11 + rank)
    ids = torch.arange(S) % SCENES_TOTAL
    data = torch.empty(S, M, 4, device=dev)
    for s0 in range(0, S, SCENES_PER_STEP):
        blk = make_batch(gen, dev, ids[s0:s0 + SCENES_PER_STEP], SCENES_TOTAL)
        data[s0:s0 + SCENES_PER_STEP] = blk.view(SCENES_PER_STEP, M, 4)
    LatentOptimizer(dec, S, num_samples=ns, seed=rank).reconstruct(data, num_iterations=2)      # warm-up (allocations)
    opt = LatentOptimizer(dec, S, num_samples=ns, seed=rank)
    barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    # ALL 800 iterations of upstream reconstruct.py: per iteration a fresh 8000-row subsample of every shape drawn on the
    # device, forward + dgrad + per-shape latent gradient + Adam; lr 5e-3 dropping to 5e-4 at iteration 400
    a.record(); hist = opt.reconstruct(data, num_iterations=iters, record_every=100); b.record()
    barrier()
    ms = max_over_ranks(a.elapsed_time(b))
    out["latent_opt"] = {"shapes_per_gpu": S, "samples_per_shape": ns, "iterations": iters, "seconds": ms / 1e3,
                         "ms_per_iteration": ms / iters, "samples_per_s": world * S * ns * iters / (ms / 1e3),
                         "mean_loss_first": float(hist[0].mean()), "mean_loss_iter_400": float(hist[4].mean()),
                         "mean_loss_last": float(hist[-1].mean()),
                         "sharding": f"shapes x{world} (replicas, no collective)"}
    del data, opt
    torch.cuda.empty_cache()
    try:
        out["config5"] = run_config5(dev, rank, world, barrier, max_over_ranks)
    except Exception as exc:  # noqa: BLE001
        out["config5"] = {"error": repr(exc)[:300]}
    return out


def run_config5(dev, rank, world, barrier, max_over_ranks):
    """BASELINE.json config #5: the experiments/double_lattice_3D network (8x512, CodeLength 2, SamplesPerScene 16000 = 125 tiles
    per scene) on 4096 synthetic scenes x 100 000 stored samples, GPU-resident (block-partitioned over the ranks), 64 scenes
    per GPU per iteration, the on-device sampler INSIDE the timed loop.  The geometry is a two-parameter analytic family
    (corner spheres of radius r1 united with axis cylinders of radius r2) standing in for the double lattice."""
    from deepsdf_b200.data import DeviceSDFSamples
    from deepsdf_b200.trainer import SceneTrainer
    total, stored, sps, spb = 4096, 100000, 16000, 64
    per_rank = total // world
    specs = dict(SPECS)
    specs.update({"CodeLength": 2, "SamplesPerScene": sps, "ScenesPerBatch": spb, "CodeRegularization": True, "CodeBound": 1.0})
    g = torch.Generator(device=dev).manual_seed(77 + rank)
    pairs, blk = [], 32
    for s0 in range(0, per_rank, blk):
        gs = torch.arange(rank * per_rank + s0, rank * per_rank + min(per_rank, s0 + blk), device=dev)
        r1 = (0.4 + 0.5 * (gs % 64).float() / 63.0).view(-1, 1)
        r2 = (0.10 + 0.3 * (gs // 64).float() / 63.0).view(-1, 1)
        xyz = torch.rand(gs.numel(), stored, 3, generator=g, device=dev) * 2.0 - 1.0
        cs = xyz.abs().amax(dim=-1) - 0.9
        for sx in (-1.0, 1.0):
            for sy in (-1.0, 1.0):
                for sz in (-1.0, 1.0):
                    c = torch.tensor([sx, sy, sz], device=dev)
                    cs = torch.maximum(cs, -((xyz - c).abs().pow(3).sum(dim=-1).pow(1.0 / 3.0) - r1))
        sdf = torch.minimum(cs, cross_ms_sdf_torch(xyz, r2))
        rows = torch.cat([xyz, sdf.unsqueeze(-1)], -1)
        for i in range(gs.numel()):
            pos = rows[i, :, 3] > 0
            pairs.append((rows[i][pos], rows[i][~pos]))
    ds = DeviceSDFSamples(None, None, sps, 3, dev, arrays=pairs)
    del pairs
    torch.cuda.empty_cache()
    tr = SceneTrainer(specs, per_rank, dev, rank=rank, world_size=world, seed=1)
    gen = torch.Generator().manual_seed(9 + rank)
    g_rows, r0 = spb * sps * world, rank * spb * sps
    tr.train_epoch(ds, 1, scenes_per_batch=spb, generator=gen, max_iterations=3, global_rows=g_rows, row0=r0)     # warm-up
    its = per_rank // spb
    epochs = max(1, -(-32 // its))
    barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    losses = []
    for e in range(epochs):
        losses += tr.train_epoch(ds, 2 + e, scenes_per_batch=spb, generator=gen, global_rows=g_rows, row0=r0)
    b.record()
    barrier()
    ms = max_over_ranks(a.elapsed_time(b))
    n_it = len(losses)
    return {"scenes_total": total, "scenes_per_gpu": per_rank, "stored_rows_per_scene": stored, "resident_GB_per_gpu": ds.total_rows * 16 / 1e9,
            "samples_per_scene": sps, "scenes_per_gpu_per_iteration": spb, "epochs_timed": epochs, "iterations_timed": n_it,
            "ms_per_iteration": ms / n_it, "samples_per_s": world * n_it * spb * sps / (ms / 1e3),
            "loss_first": float(losses[0][0]) * world, "loss_last": float(losses[-1][0]) * world,
            "what": "config #5: L=2 8x512, sampler + step per iteration, weak-scaled 64 scenes/GPU/iteration"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-graph", action="store_true", help="launch the timed steps kernel by kernel instead of replaying the captured CUDA graph")
    ap.add_argument("--no-extras", action="store_true", help="skip the grid / latent-optimisation secondary metrics")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak: 64 scenes per GPU per step (default, the driver's SCALE runs); strong: the fixed 64-scene batch split "
                         "over the GPUs.  The default line always carries a `strong` object as well")
    args = ap.parse_args()
    # stdout carries exactly one JSON line: libraries that print to fd 1 (NCCL's version banner) go to stderr instead
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(real_stdout, "w")
    if args.impl == "reference":
        run_reference(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
