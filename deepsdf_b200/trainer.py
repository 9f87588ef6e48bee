"""Fused-mode trainer: the hot loop of ``train_deep_sdf.py:472-545`` driven by ``specs.json``.

``SceneTrainer`` is what the build's own training entry point calls per iteration.  It keeps everything the reference
keeps on the host (``lat_vecs``, its Adam moments) resident in HBM, and sequences the C-ABI kernels on the current
stream:

    scene_prepare (Embedding max_norm renorm + per-scene latent bias)     train_deep_sdf.py:385,509
    fused forward  (+ clamp, L1 loss, dL/dy)                              :511-519
    fused dgrad chain, split-K wgrad, finish (weight-norm backward, latent gradient, regulariser, loss)   :523-533
    [NCCL all-reduce of the flat decoder gradient when world_size > 1]    (replaces nn.DataParallel, :353)
    [gradient clipping]                                                   :541-543
    Adam on the flat decoder parameters and on ALL latent rows (dense)    :400-411,545
    weight-norm fold + bf16 repack for the next step

Multi-GPU (one process per GPU): scenes are partitioned across ranks; every rank owns the latent rows, Adam moments and
samples of its scenes; only the 7.4 MB flat decoder gradient is exchanged.  The loss normaliser is the GLOBAL number of
rows (:519), so each rank's gradient is already scaled and the all-reduce is a plain sum.  Dropout streams are keyed by
the global row id, so results do not depend on the number of ranks.

``replicated=True`` is the strict mode of SURVEY.md 8(e): every rank keeps ALL latent rows (and the whole sample store),
batches are drawn from the reference's GLOBAL scene shuffle (:369-375) with a generator shared by all ranks, rank r steps
the r-th contiguous share of each batch, and the dense latent gradient rides in the same all-reduce as the decoder
gradient (one flat buffer), so that every rank applies the identical Adam update to every latent row.  An N-rank run then
walks the same batches, samples and dropout masks as the 1-rank run and differs only in floating-point summation order.

Python here is plumbing (tensors, streams, torch.distributed); every arithmetic step is a kernel of libdsdf_b200.so.
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Sequence

import torch

from . import _lib as L
from .engine import FusedEngine, TILE
from .networks.deep_sdf_decoder import Decoder


# ---------------------------------------------------------------------------------------------- LR schedules
class LearningRateSchedule:
    """train_deep_sdf.py:23-93: ``Step`` / ``Warmup`` / ``Constant`` entries of specs["LearningRateSchedule"]."""

    def __init__(self, spec: Dict):
        self.kind = spec["Type"]
        if self.kind == "Step":
            self.initial, self.interval, self.factor = spec["Initial"], spec["Interval"], spec["Factor"]
        elif self.kind == "Warmup":
            self.initial, self.warmed_up, self.length = spec["Initial"], spec["Final"], spec["Length"]
        elif self.kind == "Constant":
            self.value = spec["Value"]
        else:
            raise Exception('no known learning rate schedule of type "{}"'.format(self.kind))

    def get_learning_rate(self, epoch: int) -> float:
        if self.kind == "Step":
            return self.initial * (self.factor ** (epoch // self.interval))
        if self.kind == "Warmup":
            if epoch > self.length:
                return self.warmed_up
            return self.initial + (self.warmed_up - self.initial) * epoch / self.length
        return self.value


def get_spec_with_default(specs: Dict, key: str, default):
    """train_deep_sdf.py:235-239."""
    return specs[key] if key in specs else default


def _flatten_into(tensors: Sequence[torch.Tensor], align: int = 64):
    """One flat fp32 buffer holding all tensors back to back (each start aligned to ``align`` floats)."""
    offs, total = [], 0
    for t in tensors:
        offs.append(total)
        total += (t.numel() + align - 1) // align * align
    flat = torch.zeros(total, dtype=torch.float32, device=tensors[0].device)
    for t, o in zip(tensors, offs):
        flat[o:o + t.numel()].copy_(t.detach().reshape(-1))
    return flat, offs


class SceneTrainer:
    def __init__(self, specs: Dict, num_scenes: int, device, rank: int = 0, world_size: int = 1, process_group=None,
                 seed: int = 0, decoder: Optional[Decoder] = None, replicated: bool = False):
        """``specs`` is a parsed ``specs.json``; ``num_scenes`` the number of scenes THIS rank owns (``replicated``: all
        scenes of the split, identical on every rank)."""
        self.specs = specs
        self.device = torch.device(device)
        self.rank, self.world_size, self.pg = rank, world_size, process_group
        self.latent_size = specs["CodeLength"]
        if decoder is None:
            gen_state = torch.random.get_rng_state()
            torch.manual_seed(seed)                       # identical decoder init on every rank
            decoder = Decoder(self.latent_size, **specs["NetworkSpecs"])
            torch.random.set_rng_state(gen_state)
        self.decoder = decoder.to(self.device)
        self.decoder.train()
        # ---- flat parameter / gradient / Adam-moment buffers; the module's Parameters become views into ``flat_p``
        plist = self.decoder.flat_param_list()
        self.flat_p, self._offs = _flatten_into(plist)
        self.replicated = bool(replicated)
        # strict mode: decoder and latent gradients share one buffer = one all-reduce
        n_lat = num_scenes * self.latent_size if (self.replicated and world_size > 1) else 0
        self._nvls = None
        self._gbuf = self._alloc_grad_buffer(self.flat_p.numel() + n_lat)
        self.flat_g = self._gbuf[:self.flat_p.numel()]
        self.exp_avg = torch.zeros_like(self.flat_p)
        self.exp_avg_sq = torch.zeros_like(self.flat_p)
        self._grad_views: List[torch.Tensor] = []
        for p, o in zip(plist, self._offs):
            p.data = self.flat_p[o:o + p.numel()].view(p.shape)
            self._grad_views.append(self.flat_g[o:o + p.numel()].view(p.shape))
        self.engine = FusedEngine(self.decoder)
        self.PG = self.engine.grads_struct(self._grad_views)
        # ---- latent codes of the scenes this rank owns (train_deep_sdf.py:385-390)
        self.num_scenes = num_scenes
        std = get_spec_with_default(specs, "CodeInitStdDev", 1.0) / math.sqrt(self.latent_size)
        g = torch.Generator(device="cpu").manual_seed(seed * 7919 + 17 + (0 if self.replicated else rank))
        self.lat_vecs = (torch.randn(num_scenes, self.latent_size, generator=g) * std).to(self.device)
        self.lat_grad = self._gbuf[self.flat_p.numel():].view(num_scenes, self.latent_size) if n_lat else torch.zeros_like(self.lat_vecs)
        self.lat_exp_avg = torch.zeros_like(self.lat_vecs)
        self.lat_exp_avg_sq = torch.zeros_like(self.lat_vecs)
        # ---- hyper-parameters
        self.lr_schedules = [LearningRateSchedule(s) for s in specs["LearningRateSchedule"]]
        self.clamp = specs["ClampingDistance"]                                          # :300-302
        self.enforce_minmax = True
        self.code_bound = get_spec_with_default(specs, "CodeBound", None)
        self.do_code_reg = get_spec_with_default(specs, "CodeRegularization", True)
        self.code_reg_lambda = get_spec_with_default(specs, "CodeRegularizationLambda", 1e-4)
        self.grad_clip = get_spec_with_default(specs, "GradientClipNorm", None)
        self.step_count = 0
        # device-resident step state (include/dsdf.h dsdf_step_state): optimiser step counter + Adam bias corrections.  The
        # dropout stream and both Adam updates read it from memory, so a step has no per-step kernel parameter and can be
        # replayed as one CUDA graph (step_graphed)
        self.step_state = torch.zeros(16, dtype=torch.int32, device=self.device)
        self.use_graph = False        # train_epoch replays the captured step (step_graphed) instead of launching it kernel by kernel
        self.batch_split = 1          # --batch_split (train_deep_sdf.py:602-609): chunks per batch, gradients accumulated
        self.seed = seed
        self._clip_scratch = torch.zeros(1024, device=self.device)
        self._clip_out = torch.ones(2, device=self.device)
        self._tile_scene_cache: Dict = {}
        # second stream: per-scene sums, latent gradient and latent Adam run under the wgrad kernel (engine.loss_and_grads)
        # (DSDF_NO_SIDE_STREAM=1 keeps everything on one stream: A/B timing)
        import os
        self.side_stream = None if os.environ.get("DSDF_NO_SIDE_STREAM") else torch.cuda.Stream(device=self.device)

    # ------------------------------------------------------------------ helpers
    def _alloc_grad_buffer(self, numel: int) -> torch.Tensor:
        """The flat gradient buffer.  With several ranks it is placed in symmetric memory mapped through an NVSwitch multicast
        address, so that the exchange of the step can be this library's own all-reduce kernel (``dsdf_nvls_allreduce_f32``,
        csrc/nvls_allreduce.cu) instead of ncclAllReduce.  torch.distributed._symmetric_memory is used for the allocation and the
        rendezvous only.  All ranks must agree, so the outcome is MIN-reduced; any rank that cannot (no multicast support, backend
        other than NCCL, ``DSDF_ALLREDUCE=nccl``) sends everybody to the NCCL path."""
        import os
        dev = self.flat_p.device
        if self.world_size <= 1:
            return torch.zeros(numel, dtype=torch.float32, device=dev)
        import torch.distributed as dist
        group = self.pg if self.pg is not None else dist.group.WORLD
        buf, hdl, ok = None, None, 0
        if os.environ.get("DSDF_ALLREDUCE", "nvls") == "nvls" and dist.get_backend(group) == "nccl" and numel % 4 == 0:
            try:
                import torch.distributed._symmetric_memory as symm
                buf = symm.empty(numel, dtype=torch.float32, device=dev)
                buf.zero_()
                hdl = symm.rendezvous(buf, group.group_name)
                words = int(self.engine_lib().dsdf_nvls_signal_words(self.world_size))
                ok = int(hdl.multicast_ptr != 0 and hdl.signal_pad_size >= 4 * words and hdl.world_size == self.world_size)
            except Exception:  # noqa: BLE001 - symmetric memory is an optimisation; NCCL is the fallback for the whole group
                ok = 0
        flag = torch.tensor([ok], device=dev, dtype=torch.int32)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
        if int(flag.item()) == 1:
            self._nvls = hdl
            return buf
        return torch.zeros(numel, dtype=torch.float32, device=dev)

    @staticmethod
    def engine_lib():
        return L.lib()

    def tile_scene_for(self, n_scenes: int, samples_per_scene: int) -> torch.Tensor:
        """Scene-major batches (train_deep_sdf.py:496-499: indices repeated SamplesPerScene times): tile t of 128 rows
        belongs to batch slot t // (SamplesPerScene/128)."""
        key = (n_scenes, samples_per_scene)
        t = self._tile_scene_cache.get(key)
        if t is None:
            if samples_per_scene % TILE:
                raise L.DsdfError("the fused trainer needs SamplesPerScene % 128 == 0 (16384 and 16000 both are)")
            t = torch.arange(n_scenes, dtype=torch.int32, device=self.device).repeat_interleave(samples_per_scene // TILE)
            self._tile_scene_cache[key] = t.contiguous()
        return self._tile_scene_cache[key]

    def dropout_key(self) -> int:
        """Philox key of this trainer's dropout streams.  Row r of optimiser step s draws from counter row0 + (s << 40) + r
        (``dsdf_dropout_mask(key, layer, row0 + (s << 40), ...)`` reproduces the masks of step s)."""
        return (self.seed * 0x9E3779B1 + 0x51ED27) & 0x3FFFFFFFFFFFFFFF

    def set_step(self, step: int) -> None:
        """Resume: host and device step counters (checkpoint loading)."""
        self.step_count = int(step)
        with torch.cuda.device(self.device):
            L.check(self.engine.lib.dsdf_step_set(self.step_state.data_ptr(), int(step), L.stream_ptr(self.device)))

    def learning_rates(self, epoch: int):
        return [s.get_learning_rate(epoch) for s in self.lr_schedules]                  # :315-318

    # ------------------------------------------------------------------ one optimiser step
    def step(self, samples: torch.Tensor, scene_ids: torch.Tensor, epoch: int, global_rows: Optional[int] = None,
             row0: int = 0, batch_scene_ids: Optional[torch.Tensor] = None) -> torch.Tensor:
        """One iteration of train_deep_sdf.py:481-545 (``self.batch_split`` chunks, :495-535) on this rank's share of the batch.

        samples    [S * SamplesPerScene, 4] fp32 device tensor, scene-major rows of (x, y, z, sdf) - exactly what the
                   reference's DataLoader yields after ``reshape(-1, 4)`` (:483); read in place, never split or copied
        scene_ids  [S] int64 device tensor: rows of ``lat_vecs`` (local numbering), unique within the batch
        batch_scene_ids  strict mode only: the scene ids of the WHOLE batch (all ranks' shares), so that the Embedding
                   max_norm renorm (:385) touches the same rows of every replica of ``lat_vecs``
        Returns the device tensor [2] = (loss incl. regulariser, data term) of this rank's rows, normalised by the
        global row count like the reference's ``chunk_loss`` (:519).
        """
        eng = self.engine
        S = scene_ids.numel()
        N = samples.shape[0]
        sps = N // S
        n_total = int(global_rows) if global_rows is not None else N * self.world_size
        tile_scene = self.tile_scene_for(S, sps)
        lam = 0.0
        if self.do_code_reg:
            lam = self.code_reg_lambda * min(1.0, epoch / 100.0)                        # :523-529
        self.step_count += 1
        with torch.cuda.device(self.device):
            L.check(eng.lib.dsdf_step_advance(self.step_state.data_ptr(), 0.9, 0.999, L.stream_ptr(self.device)))
        seed = self.dropout_key()               # constant Philox key; the step enters through the row counter (step << 40)
        flat = samples.view(-1)
        code_bound = self.code_bound
        shared_lat = self.replicated and self.world_size > 1
        if shared_lat:
            if batch_scene_ids is None:
                raise L.DsdfError("replicated latents: step() needs the whole batch's scene ids (batch_scene_ids)")
            if code_bound:
                eng.scene_bias(self.lat_vecs, batch_scene_ids, code_bound)     # renorm only; the bias it writes is scratch
                code_bound = None
        clamp = self.clamp if self.enforce_minmax else 0.0
        n_tiles = N // TILE
        k_split = max(1, min(int(self.batch_split), n_tiles))
        lr_dec, lr_lat = self.learning_rates(epoch)
        lat_adam_done = False
        if k_split == 1:
            # (the latent Adam must NOT ride the side stream: the parameter phase of finish still reads the pre-update codes
            #  for the gradients of the latent weight columns; it runs after finish, beside the all-reduce)
            loss = eng.loss_and_grads(self.lat_vecs, scene_ids, tile_scene, flat, flat[3:], clamp=clamp,
                                      reg_lambda_eff=lam, n_total=n_total, code_bound=code_bound, seed=seed,
                                      row0=row0, PG=self.PG, lat_grad=self.lat_grad, training=True, point_stride=4, gt_stride=4,
                                      n_rows=N, side_stream=self.side_stream, step_state=self.step_state)
        else:
            # train_deep_sdf.py:495-535: the batch in `batch_split` row chunks (torch.chunk sizes: ceil(rows / k), here in whole
            # 128-row tiles), every chunk's loss divided by the WHOLE batch's row count (:519), gradients accumulated, one
            # optimiser step.  Dropout streams are keyed by the global row id, so chunking leaves the masks untouched; the
            # activation stash is sized by the largest chunk only.
            tps = sps // TILE
            per = (n_tiles + k_split - 1) // k_split
            for c in range(k_split):
                t0, t1 = c * per, min(n_tiles, (c + 1) * per)
                if t0 >= t1:
                    break
                s0, s1 = t0 // tps, (t1 - 1) // tps + 1               # scenes this chunk touches (a scene may span chunks)
                key = ("chunk", S, sps, k_split, c)
                ts = self._tile_scene_cache.get(key)
                if ts is None:
                    ts = (torch.arange(t0, t1, dtype=torch.int32, device=self.device) // tps - s0).contiguous()
                    self._tile_scene_cache[key] = ts
                part = flat[t0 * TILE * 4:]
                loss = eng.loss_and_grads(self.lat_vecs, scene_ids[s0:s1], ts, part, part[3:], clamp=clamp, reg_lambda_eff=lam,
                                          n_total=n_total, code_bound=code_bound, seed=seed, row0=row0 + t0 * TILE, PG=self.PG,
                                          lat_grad=self.lat_grad, training=True, point_stride=4, gt_stride=4,
                                          n_rows=(t1 - t0) * TILE, accumulate=c > 0, scene_capacity=S, step_state=self.step_state)
        # the one exchange of the step: all-reduce (sum) of the flat decoder gradient (strict mode: + the dense latent gradient
        # behind it in the same buffer).  Replaces nn.DataParallel's gradient reduction (train_deep_sdf.py:353).
        work = None
        ev = getattr(self, "comm_events", None)      # optional timing hooks (tools/allreduce_exposure.py): (issued, joined) pairs
        if ev is not None:
            e_issue = torch.cuda.Event(enable_timing=True)
            e_issue.record()
        if self._nvls is not None:
            # this library's own exchange: multimem.ld_reduce / multimem.st over the NVSwitch multicast mapping of the buffer
            # (stream-ordered behind finish; nothing to overlap it with but the latent Adam, which is 8 us)
            with torch.cuda.device(self.device):
                L.check(eng.lib.dsdf_nvls_allreduce_f32(self._nvls.multicast_ptr, self._nvls.signal_pad_ptrs_dev, self._nvls.rank,
                                                        self.world_size, self._gbuf.numel(), L.stream_ptr(self.device)))
        elif self.world_size > 1:
            import torch.distributed as dist
            work = dist.all_reduce(self._gbuf, op=dist.ReduceOp.SUM, group=self.pg, async_op=True)
        if not lat_adam_done:
            if shared_lat and work is not None:          # the latent gradient is part of the exchange
                work.wait()
                work = None
            eng.adam(self.lat_vecs, self.lat_grad, self.lat_exp_avg, self.lat_exp_avg_sq, self.step_count, lr_lat, step_state=self.step_state)
        if work is not None:
            work.wait()
        if ev is not None:
            e_join = torch.cuda.Event(enable_timing=True)
            e_join.record()
            ev.append((e_issue, e_join))
        scale_dev = None
        if self.grad_clip is not None:                                                  # :541-543 (decoder only)
            eng.clip_coef(self.flat_g, self.grad_clip, self._clip_scratch, self._clip_out)
            scale_dev = self._clip_out[1:2]
        eng.adam(self.flat_p, self.flat_g, self.exp_avg, self.exp_avg_sq, self.step_count, lr_dec, scale_dev=scale_dev,
                 step_state=self.step_state)
        # the kernel wrote the parameters behind torch's version counters: cached packs of other engines key on this
        object.__setattr__(self.decoder, "_param_epoch", getattr(self.decoder, "_param_epoch", 0) + 1)
        eng.repack(train_only=True)      # the training variants only; an eval-mode call on this engine rebuilds the others
        return loss

    # ------------------------------------------------------------------ the same step as ONE CUDA graph
    def step_graphed(self, samples: torch.Tensor, scene_ids: torch.Tensor, epoch: int, global_rows: Optional[int] = None,
                     row0: int = 0) -> torch.Tensor:
        """``step`` replayed from a captured CUDA graph (partition mode, batch_split 1): ~20 kernel launches, the side-stream
        fork / join and the NCCL all-reduce become one graph launch, which removes the launch gaps that matter once a rank's
        share of the batch is small (strong scaling).  Nothing in the step depends on a per-step kernel PARAMETER (dropout
        stream and Adam bias corrections read the device step state), so a replay is a full optimiser step.  The graph is
        re-captured when the shapes or an epoch-dependent scalar (learning rates, regulariser ramp) change.  ``samples`` /
        ``scene_ids`` are copied into the graph's static input buffers (or ARE them: ``graph_inputs``).  Returns the static
        loss tensor [2] (overwritten by the next replay)."""
        if (self.replicated and self.world_size > 1) or self.batch_split != 1:
            return self.step(samples, scene_ids, epoch, global_rows=global_rows, row0=row0)
        S, N = scene_ids.numel(), samples.shape[0]
        lam = self.code_reg_lambda * min(1.0, epoch / 100.0) if self.do_code_reg else 0.0
        key = (N, S, global_rows, row0, lam, tuple(self.learning_rates(epoch)))
        if getattr(self, "_graph_key", None) != key:
            self._capture(key, samples, scene_ids, epoch, global_rows, row0)
        g_s, g_i = self._graph_in
        if samples.data_ptr() != g_s.data_ptr():
            g_s.copy_(samples, non_blocking=True)
        if scene_ids.data_ptr() != g_i.data_ptr():
            g_i.copy_(scene_ids, non_blocking=True)
        self._graph.replay()
        self.step_count += 1
        object.__setattr__(self.decoder, "_param_epoch", getattr(self.decoder, "_param_epoch", 0) + 1)
        return self._graph_loss

    def release_graph(self) -> None:
        """Drop the captured graph.  MUST run before ``torch.distributed.destroy_process_group()`` when world_size > 1: a
        live CUDA graph that holds NCCL kernels keeps the communicator busy and its teardown never returns."""
        torch.cuda.synchronize(self.device)
        self._graph = None
        self._graph_loss = None
        self._graph_key = None
        import gc
        gc.collect()
        torch.cuda.synchronize(self.device)

    def graph_inputs(self, n_rows: int, n_scenes: int):
        """Static input buffers (samples [n_rows, 4], scene_ids [n_scenes]) a caller may fill directly (e.g. the H2D copy of
        the next batch) before ``step_graphed`` - no device-to-device copy then."""
        gi = getattr(self, "_graph_in", None)
        if gi is None or gi[0].shape[0] != n_rows or gi[1].numel() != n_scenes:
            self._graph_in = (torch.empty(n_rows, 4, device=self.device), torch.zeros(n_scenes, dtype=torch.int64, device=self.device))
            self._graph_key = None
        return self._graph_in

    def _capture(self, key, samples, scene_ids, epoch, global_rows, row0) -> None:
        g_s, g_i = self.graph_inputs(samples.shape[0], scene_ids.numel())
        g_s.copy_(samples)
        g_i.copy_(scene_ids)
        # warm-up step on the real inputs (lazy allocations, kernel attributes, NCCL communicator), then put the state back
        saved = [t.clone() for t in (self.flat_p, self.exp_avg, self.exp_avg_sq, self.lat_vecs, self.lat_exp_avg, self.lat_exp_avg_sq,
                                     self.step_state)]
        count = self.step_count
        cap_stream = torch.cuda.Stream(device=self.device)
        cap_stream.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(cap_stream):
            self.step(g_s, g_i, epoch, global_rows=global_rows, row0=row0)
        torch.cuda.current_stream(self.device).wait_stream(cap_stream)
        torch.cuda.synchronize(self.device)
        for t, s in zip((self.flat_p, self.exp_avg, self.exp_avg_sq, self.lat_vecs, self.lat_exp_avg, self.lat_exp_avg_sq, self.step_state),
                        saved):
            t.copy_(s)
        self.step_count = count
        self.engine.repack(train_only=True)
        n0 = int(self.engine.lib.dsdf_launch_count())
        self._graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self._graph):
            self._graph_loss = self.step(g_s, g_i, epoch, global_rows=global_rows, row0=row0)
        self.graph_launches = int(self.engine.lib.dsdf_launch_count()) - n0      # this repo's kernels inside one replay
        self.step_count = count                                                  # capture executes nothing
        self._graph_key = key

    def step_from_host(self, samples_host: torch.Tensor, scene_ids_host: torch.Tensor, epoch: int, staging: Dict,
                       global_rows: Optional[int] = None, row0: int = 0, graphed: bool = False) -> torch.Tensor:
        """The call a user of the reference's DataLoader makes: pinned host samples in, device loss out.  ``staging`` holds
        the reusable device buffers ("samples", "scene_ids"); ``graphed``: the H2D copies land in the captured step's static
        input buffers and the step is one graph replay (``step_graphed``)."""
        if graphed:
            d_s, d_i = self.graph_inputs(samples_host.shape[0], scene_ids_host.numel())
        else:
            d_s, d_i = staging["samples"], staging["scene_ids"]
        d_s.copy_(samples_host, non_blocking=True)
        d_i.copy_(scene_ids_host, non_blocking=True)
        return (self.step_graphed if graphed else self.step)(d_s, d_i, epoch, global_rows=global_rows, row0=row0)

    def train_epoch(self, dataset, epoch: int, scenes_per_batch: Optional[int] = None, generator=None,
                    max_iterations: Optional[int] = None, global_rows: Optional[int] = None, row0: int = 0) -> List[torch.Tensor]:
        """One epoch of train_deep_sdf.py:472-545 from a ``DeviceSDFSamples`` store: shuffled scene batches (drop_last),
        every batch subsampled on the device (``dsdf_sample_batch``) and stepped.  Returns the per-iteration loss tensors
        (device, no synchronisation - the reference's ``.item()`` per chunk, :535, is the caller's choice).  Multi-GPU:
        ``scenes_per_batch`` is this rank's share, ``global_rows`` the whole batch's rows, ``row0`` this rank's offset.
        Strict mode (``replicated``): ``dataset`` holds every scene, ``generator`` is seeded identically on all ranks,
        ``scenes_per_batch`` is the GLOBAL batch; the share, ``global_rows`` and ``row0`` follow from the rank."""
        from .data import epoch_batches
        spb = scenes_per_batch or self.specs["ScenesPerBatch"]
        shared_lat = self.replicated and self.world_size > 1
        if shared_lat and spb % self.world_size:
            raise L.DsdfError("replicated latents: scenes_per_batch must be a multiple of the number of ranks")
        share = spb // self.world_size if shared_lat else spb
        losses = []
        buf = None
        for it, ids in enumerate(epoch_batches(len(dataset), spb, generator)):
            if max_iterations is not None and it >= max_iterations:
                break
            ids_d = ids.to(self.device)
            mine = ids_d[self.rank * share:(self.rank + 1) * share].contiguous() if shared_lat else ids_d
            seed = (self.seed * 0x51ED27 + epoch * 0x10001 + it) & 0x7FFFFFFFFFFFFFFF
            buf = dataset.sample_batch(mine, seed, out=buf)
            if shared_lat:
                sps = buf.shape[0] // share
                loss = self.step(buf, mine, epoch, global_rows=spb * sps, row0=self.rank * share * sps, batch_scene_ids=ids_d)
            else:
                loss = (self.step_graphed if self.use_graph else self.step)(buf, mine, epoch, global_rows=global_rows, row0=row0)
            losses.append(loss.clone())
        return losses

    # ------------------------------------------------------------------ checkpoint surface (train_deep_sdf.py:96-143)
    def latent_state_dict(self) -> Dict[str, torch.Tensor]:
        return {"weight": self.lat_vecs.detach().cpu().clone()}

    def mean_latent_magnitude(self) -> float:
        return float(torch.mean(torch.norm(self.lat_vecs.detach(), dim=1)))             # :242-243


class HostFeeder:
    """Double-buffered host -> device input pipeline in front of ``SceneTrainer`` - the part the reference's DataLoader
    workers + ``.cuda()`` play (train_deep_sdf.py:369-375, 483-509).  ``submit`` starts the copy of a pinned host batch on a
    copy stream; ``step`` trains on the oldest submitted batch.  Submitting batch i+1 before stepping batch i puts its copy
    under that step's kernels, so a host-fed loop runs at the device-resident step time instead of copy + step."""

    def __init__(self, trainer: "SceneTrainer", n_rows: int, n_scenes: int, depth: int = 2, graphed: bool = True):
        self.trainer, self.depth, self.graphed = trainer, depth, graphed
        dev = trainer.device
        self.copy_stream = torch.cuda.Stream(device=dev)
        self._bufs = [(torch.empty(n_rows, 4, device=dev), torch.zeros(n_scenes, dtype=torch.int64, device=dev)) for _ in range(depth)]
        self._ready = [torch.cuda.Event() for _ in range(depth)]
        self._consumed = [torch.cuda.Event() for _ in range(depth)]
        self._head = self._tail = 0

    def submit(self, samples_host: torch.Tensor, scene_ids_host: torch.Tensor) -> None:
        if self._head - self._tail >= self.depth:
            raise L.DsdfError("HostFeeder: all staging buffers are in flight; call step() first")
        k = self._head % self.depth
        d_s, d_i = self._bufs[k]
        if self._head >= self.depth:
            self.copy_stream.wait_event(self._consumed[k])     # the step that read this buffer last has been enqueued and must finish
        with torch.cuda.stream(self.copy_stream):
            d_s.copy_(samples_host, non_blocking=True)
            d_i.copy_(scene_ids_host, non_blocking=True)
            self._ready[k].record(self.copy_stream)
        self._head += 1

    def step(self, epoch: int, global_rows: Optional[int] = None, row0: int = 0) -> torch.Tensor:
        if self._tail >= self._head:
            raise L.DsdfError("HostFeeder: nothing submitted")
        k = self._tail % self.depth
        d_s, d_i = self._bufs[k]
        cur = torch.cuda.current_stream(self.trainer.device)
        cur.wait_event(self._ready[k])
        fn = self.trainer.step_graphed if self.graphed else self.trainer.step
        loss = fn(d_s, d_i, epoch, global_rows=global_rows, row0=row0)
        self._consumed[k].record(cur)
        self._tail += 1
        return loss

    def step_async(self, epoch: int, global_rows: Optional[int] = None, row0: int = 0) -> "LossHandle":
        """``step`` whose loss is read back asynchronously: the [2] loss tensor is copied into a pinned ring slot behind the
        step, and the returned handle's ``wait()`` blocks until THAT step has finished.  Reading the previous step's handle
        after queueing the next step keeps the GPU busy across the read (the reference's ``.item()`` per iteration,
        train_deep_sdf.py:535, drains the queue every step).  A handle must be waited for before ``depth`` further steps reuse
        its slot."""
        if not hasattr(self, "_loss_host"):
            self._loss_host = [torch.empty(2, dtype=torch.float32).pin_memory() for _ in range(self.depth)]
            self._loss_done = [torch.cuda.Event() for _ in range(self.depth)]
        k = self._tail % self.depth
        loss = self.step(epoch, global_rows=global_rows, row0=row0)
        self._loss_host[k].copy_(loss, non_blocking=True)
        self._loss_done[k].record(torch.cuda.current_stream(self.trainer.device))
        return LossHandle(self._loss_host[k], self._loss_done[k])


class LossHandle:
    """Result of ``HostFeeder.step_async``: ``wait()`` returns the step's (total, data-term) loss as a CPU tensor."""

    def __init__(self, host: torch.Tensor, done: torch.cuda.Event):
        self._host, self._done = host, done

    def wait(self) -> torch.Tensor:
        self._done.synchronize()
        return self._host.clone()

