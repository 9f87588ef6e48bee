"""Host-side driver of the fused (tcgen05) scene-mode path.

``FusedEngine`` wraps a drop-in ``Decoder`` whose NetworkSpecs the fused kernels support (8x512 class) and exposes
the calls the build's trainer / grid evaluator / latent optimiser use:

  * ``repack()``                      fold weight-norm + bf16 pack (once per parameter update)
  * ``decode_scenes(...)``            SDF of points grouped in 128-row scene-aligned tiles (eval or train mode)
  * ``loss_and_grads(...)``           one training step's loss and gradients (train_deep_sdf.py:483-533)

All compute is in the C-ABI library; this file only allocates tensors and sequences the calls on the current stream.
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, List, Optional, Tuple

import torch

from . import _lib as L
from .networks.deep_sdf_decoder import Decoder

TILE = 128
STORE_DYS_TRAIN = 0xFFFFFFFE      # include/dsdf.h DSDF_STORE_DYS_TRAIN
SUMS_ALL = 0xFFFFFFFF


class TrainBufs(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ("acts", "dys", "bitmask", "dy", "tile_loss", "tile_part", "wg_part", "last_part",
                                           "scene_sums")]


def _declare_train(lib) -> None:
    if getattr(lib, "_dsdf_train_declared", False):
        return
    vp, i32, i64, f32, sz = C.c_void_p, C.c_int32, C.c_int64, C.c_float, C.c_size_t
    P = C.POINTER
    lib.dsdf_train_buf_sizes.restype = C.c_int
    lib.dsdf_train_buf_sizes.argtypes = [P(L.NetDesc), i64, i32, P(sz)]
    lib.dsdf_fused_dgrad.restype = C.c_int
    lib.dsdf_fused_dgrad.argtypes = [P(L.NetDesc), vp, P(TrainBufs), vp, i32, i64, i32, C.c_uint32, C.c_uint32, vp]
    lib.dsdf_fused_input_grad.restype = C.c_int
    lib.dsdf_fused_input_grad.argtypes = [P(L.NetDesc), vp, P(TrainBufs), i64, vp, i64, vp]
    lib.dsdf_fused_latent_finish.restype = C.c_int
    lib.dsdf_fused_latent_finish.argtypes = [P(L.NetDesc), vp, P(TrainBufs), vp, i64, vp, vp, i32, f32, f32, vp, vp, vp]
    lib.dsdf_fused_wgrad.restype = C.c_int
    lib.dsdf_fused_wgrad.argtypes = [P(L.NetDesc), P(TrainBufs), i64, vp]
    lib.dsdf_fused_finish.restype = C.c_int
    lib.dsdf_fused_finish.argtypes = [P(L.NetDesc), vp, P(L.Params), P(TrainBufs), vp, i64, vp, vp, i32, i64, f32, f32, i32, i32, i32,
                                      P(L.ParamGrads), vp, vp, vp]
    lib.dsdf_adam.restype = C.c_int
    lib.dsdf_adam.argtypes = [vp, vp, vp, vp, i64, f32, f32, f32, f32, i32, f32, vp, vp]
    lib.dsdf_clip_coef.restype = C.c_int
    lib.dsdf_clip_coef.argtypes = [vp, i64, f32, vp, vp, vp]
    lib._dsdf_train_declared = True


class FusedEngine:
    def __init__(self, decoder: Decoder):
        self.lib = L.lib()
        _declare_train(self.lib)
        self.decoder = decoder
        self.spec = decoder.spec
        self.desc = decoder._desc
        if self.lib.dsdf_fused_supported(C.byref(self.desc)) != 1:
            raise L.DsdfError("this NetworkSpecs is not supported by the fused tcgen05 path (needs 8x512-class dims); "
                              "use Decoder.forward (generic path)")
        p0 = next(decoder.parameters())
        L.require_cuda(p0)
        self.device = p0.device
        with torch.cuda.device(self.device):
            L.check(self.lib.dsdf_check_device())
        nbytes = self.lib.dsdf_pack_bytes(C.byref(self.desc))
        raw = torch.zeros(nbytes + 1024, dtype=torch.uint8, device=self.device)
        off = (-raw.data_ptr()) % 1024
        self._pack_raw = raw
        self.pack = raw[off:off + nbytes]
        self.num_hidden = self.spec.num_linear - 1
        self._bufs_key = None
        self._bufs: Dict[str, torch.Tensor] = {}
        self.kernel_events = None      # set to a list to collect (name, start_event, end_event) per fused kernel call
        self._eval_fresh = False
        self.repack()

    # ------------------------------------------------------------------ weights
    def _st(self) -> int:
        return L.stream_ptr(self.device)

    def repack(self, force: bool = True, train_only: bool = False) -> None:
        """Fold g*v/||v|| and rebuild the bf16 operand packs from the decoder's *current* parameters.  ``force=False``
        skips the launch when no parameter tensor changed since the last pack (``small.param_key``: torch's version
        counters + the trainer's ``_param_epoch``).  ``train_only`` rewrites only the training variants of the big blocks (the
        dropout scale folded into the next layer's weights): what a training loop needs every step; the eval variants are
        then rebuilt on demand by the first eval-mode call (``_ensure_eval``)."""
        from .small import param_key
        key = param_key(self.decoder)
        if not force and key == getattr(self, "_pack_key", None) and (train_only or self._eval_fresh):
            return
        self._P = self.decoder.c_params()
        PACK_EVAL, PACK_TRAIN, PACK_ZERO = 1, 2, 4
        flags = (PACK_TRAIN if train_only else PACK_EVAL | PACK_TRAIN) | (0 if getattr(self, "_pack_key", None) is not None else PACK_ZERO)
        with torch.cuda.device(self.device):
            L.check(self.lib.dsdf_pack_weights_ex(C.byref(self.desc), C.byref(self._P), self.pack.data_ptr(), flags, self._st()))
        self._pack_key = key
        self._eval_fresh = not train_only

    def _ensure_eval(self) -> None:
        """Eval-mode entry points read the plain (unscaled) blocks: rebuild them if the last re-pack skipped them."""
        if not self._eval_fresh:
            self.repack(force=True, train_only=False)

    def _b_last(self) -> torch.Tensor:
        return getattr(self.decoder, "lin" + str(self.spec.num_linear - 1)).bias

    # ------------------------------------------------------------------ scenes
    def scene_bias(self, lat_table: torch.Tensor, scene_ids: torch.Tensor, max_norm: Optional[float] = None,
                   out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """Per-scene aux rows [S, 2, 512, 16] bf16 (bias + latent part of lin0 / the latent_in layer and the xyz columns, in the
        K-column order of the forward's aux tensor-core step); applies nn.Embedding(max_norm) renorm to the listed rows in
        place."""
        L.require_cuda(lat_table, scene_ids)
        assert lat_table.dtype == torch.float32 and lat_table.is_contiguous() and lat_table.shape[1] == self.spec.latent_size
        assert scene_ids.dtype == torch.int64 and scene_ids.is_contiguous()
        S = scene_ids.numel()
        cb = out if out is not None else torch.empty(S, 2, 512, 16, dtype=torch.bfloat16, device=self.device)
        assert cb.dtype == torch.bfloat16 and cb.numel() == S * 2 * 512 * 16 and cb.is_contiguous()
        with torch.cuda.device(self.device):
            L.check(self.lib.dsdf_scene_prepare(C.byref(self.desc), self.pack.data_ptr(), lat_table.data_ptr(),
                                                scene_ids.data_ptr(), S, float(max_norm or 0.0), cb.data_ptr(), self._st()))
        return cb

    def _exact_pack(self) -> torch.Tensor:
        """(hi, lo) bf16 split of the folded hidden weights for ``decode_scenes(exact=True)``; rebuilt when the parameters changed."""
        from .small import param_key
        key = param_key(self.decoder)
        if getattr(self, "_xpack_key", None) != key:
            if getattr(self, "_xpack", None) is None:
                self._xpack = torch.empty(self.lib.dsdf_exact_pack_bytes(C.byref(self.desc)) // 2, dtype=torch.bfloat16, device=self.device)
            P = self.decoder.c_params()
            with torch.cuda.device(self.device):
                L.check(self.lib.dsdf_exact_pack(C.byref(self.desc), C.byref(P), self._xpack.data_ptr(), self._st()))
            self._xpack_key = key
        return self._xpack

    def scene_bias_exact(self, lat_table: torch.Tensor, scene_ids: torch.Tensor) -> torch.Tensor:
        """Per-scene aux rows for ``decode_scenes(exact=True)``: the fp16 (hi, lo) counterpart of ``scene_bias`` (no renorm)."""
        L.require_cuda(lat_table, scene_ids)
        S = scene_ids.numel()
        self._ensure_eval()
        out = torch.empty(S, 2, 512, 16, dtype=torch.float16, device=self.device)
        with torch.cuda.device(self.device):
            L.check(self.lib.dsdf_exact_scene_aux(C.byref(self.desc), self.pack.data_ptr(), lat_table.data_ptr(), scene_ids.data_ptr(), S,
                                                  out.data_ptr(), self._st()))
        return out

    def decode_scenes(self, cbias: torch.Tensor, tile_scene: torch.Tensor, xyz: torch.Tensor,
                      out: Optional[torch.Tensor] = None, exact: bool = False) -> torch.Tensor:
        """Eval-mode SDF [N] of xyz [N,3]; rows [128*t, 128*t+128) use scene row ``tile_scene[t]`` of ``cbias``.
        ``exact``: the split-fp16 tensor-core mode (csrc/chain_exact.cu; ``cbias`` from ``scene_bias_exact``): <= 1e-5 against
        fp32 arithmetic instead of ~1e-3, at about 4x the time."""
        L.require_cuda(cbias, tile_scene, xyz)
        N = xyz.shape[0]
        assert xyz.dtype == torch.float32 and xyz.is_contiguous() and xyz.shape[1] == 3
        assert tile_scene.dtype == torch.int32 and tile_scene.numel() == (N + TILE - 1) // TILE
        sdf = out if out is not None else torch.empty(N, device=self.device)
        self._ensure_eval()
        if exact:
            if cbias.dtype != torch.float16:
                raise L.DsdfError("decode_scenes(exact=True) needs the aux rows of scene_bias_exact (fp16)")
            xp = self._exact_pack()
            with torch.cuda.device(self.device):
                L.check(self.lib.dsdf_fused_forward_exact(C.byref(self.desc), self.pack.data_ptr(), xp.data_ptr(), cbias.data_ptr(),
                                                          tile_scene.data_ptr(), xyz.data_ptr(), 3, N, self._b_last().data_ptr(),
                                                          sdf.data_ptr(), self._st()))
            return sdf
        io = L.TrainIO()
        io.b_last = self._b_last().data_ptr()
        io.training = 0
        with torch.cuda.device(self.device):
            L.check(self.lib.dsdf_fused_forward(C.byref(self.desc), self.pack.data_ptr(), cbias.data_ptr(), tile_scene.data_ptr(),
                                                xyz.data_ptr(), 3, N, sdf.data_ptr(), C.byref(io), self._st()))
        return sdf

    # ------------------------------------------------------------------ input gradient (mesh.py:405-454 consumers)
    @property
    def lat_layer(self) -> int:
        li = [l for l in (self.spec.latent_in or ()) if 0 < l < self.num_hidden]
        return li[0] if li else -1

    def grad_buffers(self, n_rows: int, need_dys: bool = True) -> Dict[str, torch.Tensor]:
        """Scratch of the input-gradient / latent-optimisation passes for ``n_rows`` (multiple of 128) rows: gate bits,
        dL/dy seed, the bf16 dY stash (only lin0's and the latent_in layer's regions are written) and per-tile sums."""
        key = ("grad", n_rows, need_dys)
        if self._bufs_key == key:
            return self._bufs
        nh = self.num_hidden
        n_tiles = n_rows // TILE
        self._bufs = {}                      # drop the previous set before allocating the next
        self._bufs_key = None
        b = {
            "bitmask": torch.empty(nh * n_rows * 16, dtype=torch.int32, device=self.device),
            "dy": torch.empty(n_rows, device=self.device),
            "dys": torch.empty(nh * n_rows * 512 if need_dys else 8, dtype=torch.bfloat16, device=self.device),
            "tile_loss": torch.empty(n_tiles, device=self.device),
            "tile_part": torch.empty(n_tiles * (nh + 6) * 512, device=self.device),
            "sdf": torch.empty(n_rows, device=self.device),
        }
        tb = TrainBufs()
        for n in ("bitmask", "dy", "dys", "tile_loss", "tile_part"):
            if n != "dys" or need_dys:
                setattr(tb, n, b[n].data_ptr())
        self._bufs, self._tb, self._bufs_key = b, tb, key
        return b

    def sdf_and_input_grad(self, cbias: torch.Tensor, tile_scene: torch.Tensor, xyz: torch.Tensor,
                           G: Optional[torch.Tensor] = None, sdf: Optional[torch.Tensor] = None):
        """Eval-mode SDF [N] and G [N, L+3] = dSDF/d[latent, xyz] (column order of the decoder input) of scene-tiled
        points ``xyz`` [N,3], N % 128 == 0.  Three launches: forward (records relu gates, seeds dy = 1 - tanh^2),
        dgrad chain (stashes dY of lin0 and of the latent_in layer), input-gradient GEMM."""
        L.require_cuda(cbias, tile_scene, xyz)
        N = xyz.shape[0]
        if N % TILE or N == 0:
            raise L.DsdfError("sdf_and_input_grad needs a positive multiple of 128 rows (pad the queries)")
        assert xyz.dtype == torch.float32 and xyz.is_contiguous() and xyz.shape[1] == 3
        assert tile_scene.dtype == torch.int32 and tile_scene.numel() == N // TILE
        Lz = self.spec.latent_size
        self._ensure_eval()
        b = self.grad_buffers(N)
        sdf = sdf if sdf is not None else torch.empty(N, device=self.device)
        G = G if G is not None else torch.empty(N, Lz + 3, device=self.device)
        assert G.dtype == torch.float32 and G.shape[0] == N and G.shape[1] >= Lz + 3 and G.stride(1) == 1
        io = L.TrainIO()
        io.b_last = self._b_last().data_ptr()
        io.training, io.dropout_off, io.dy_unit = 1, 1, 1
        io.dy_out = b["dy"].data_ptr()
        io.bitmask = b["bitmask"].data_ptr()
        io.n_rows_pad = N
        lat = self.lat_layer
        store = 1 | ((1 << lat) if lat > 0 else 0)
        lib, desc, st = self.lib, self.desc, self._st()
        with torch.cuda.device(self.device):
            L.check(lib.dsdf_fused_forward(C.byref(desc), self.pack.data_ptr(), cbias.data_ptr(), tile_scene.data_ptr(),
                                           xyz.data_ptr(), 3, N, sdf.data_ptr(), C.byref(io), st))
            L.check(lib.dsdf_fused_dgrad(C.byref(desc), self.pack.data_ptr(), C.byref(self._tb), xyz.data_ptr(), 3, N, 0,
                                         store, 0, st))
            L.check(lib.dsdf_fused_input_grad(C.byref(desc), self.pack.data_ptr(), C.byref(self._tb), N, G.data_ptr(),
                                              G.stride(0), st))
        return sdf, G

    def sdf_and_input_grad_rows(self, queries: torch.Tensor, G: Optional[torch.Tensor] = None, sdf: Optional[torch.Tensor] = None,
                                with_grad: bool = True):
        """Eval-mode SDF [N] (and G [N, L+3] = dSDF/d[latent, xyz]) of ``queries`` [N, L+3] whose rows carry their OWN latent
        code (``decode_sdf(decoder, None, queries)``, deep_sdf/utils.py:62-63 / mesh.py:490-497), N % 128 == 0; supported code lengths: ``dsdf_fused_rows_supported`` (L <= 16,
        or long codes such as L = 256).
        One ``G`` replaces the per-(latent, control point) jvp loop of mesh.py:405-422: the jvp along a control point is
        ``G[:, i_lat] * basis[:, i_cpt]``."""
        L.require_cuda(queries)
        N, width = queries.shape
        Lz = self.spec.latent_size
        if N == 0 or (with_grad and N % TILE):
            raise L.DsdfError("sdf_and_input_grad_rows needs a positive multiple of 128 rows (pad the queries)")
        assert queries.dtype == torch.float32 and queries.is_contiguous() and width == Lz + 3
        self._ensure_eval()
        sdf = sdf if sdf is not None else torch.empty(N, device=self.device)
        io = L.TrainIO()
        io.b_last = self._b_last().data_ptr()
        lib, desc, st = self.lib, self.desc, self._st()
        if not with_grad:
            with torch.cuda.device(self.device):
                L.check(lib.dsdf_fused_forward_rows(C.byref(desc), self.pack.data_ptr(), queries.data_ptr(), width, N, sdf.data_ptr(),
                                                    C.byref(io), st))
            return sdf, None
        b = self.grad_buffers(N)
        G = G if G is not None else torch.empty(N, Lz + 3, device=self.device)
        assert G.dtype == torch.float32 and G.shape[0] == N and G.shape[1] >= Lz + 3 and G.stride(1) == 1
        io.training, io.dropout_off, io.dy_unit = 1, 1, 1
        io.dy_out = b["dy"].data_ptr()
        io.bitmask = b["bitmask"].data_ptr()
        io.n_rows_pad = N
        lat = self.lat_layer
        store = 1 | ((1 << lat) if lat > 0 else 0)
        xyz_ptr = queries.data_ptr() + 4 * Lz
        with torch.cuda.device(self.device):
            L.check(lib.dsdf_fused_forward_rows(C.byref(desc), self.pack.data_ptr(), queries.data_ptr(), width, N, sdf.data_ptr(),
                                                C.byref(io), st))
            L.check(lib.dsdf_fused_dgrad(C.byref(desc), self.pack.data_ptr(), C.byref(self._tb), xyz_ptr, width, N, 0, store, 0, st))
            L.check(lib.dsdf_fused_input_grad(C.byref(desc), self.pack.data_ptr(), C.byref(self._tb), N, G.data_ptr(), G.stride(0), st))
        return sdf, G

    def grid_points(self, N: int, i0: int, n: int, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """Rows [i0, i0+n) of create_mesh's N^3 sample grid (deep_sdf/mesh.py:37-53) as [n,3] fp32."""
        out = out if out is not None else torch.empty(n, 3, device=self.device)
        with torch.cuda.device(self.device):
            L.check(self.lib.dsdf_grid_points(N, i0, n, out.data_ptr(), self._st()))
        return out

    # ------------------------------------------------------------------ frozen-decoder latent optimisation
    def latent_loss_and_grad(self, lat_table: torch.Tensor, scene_ids: torch.Tensor, tile_scene: torch.Tensor,
                             tile_rows: Optional[torch.Tensor], samples: torch.Tensor, *, clamp: float, rows_per_shape: int,
                             l2reg: float, lat_grad: torch.Tensor, shape_loss: torch.Tensor) -> None:
        """One iteration of upstream reconstruct.py for every shape of the batch: ``samples`` [S * T * 128, 4] (x,y,z,sdf),
        shape-major, each shape padded to T tiles with ``tile_rows`` giving the valid rows of every tile.  Writes
        lat_grad [S, L] = d(mean|clamp(pred)-clamp(gt)| + l2reg*mean(z^2))/dz and shape_loss [S]."""
        N = samples.shape[0]
        S = scene_ids.numel()
        self._ensure_eval()
        b = self.grad_buffers(N, need_dys=False)
        if "cbias" not in b or b["cbias"].shape[0] != S:
            b["cbias"] = torch.empty(S, 2, 512, 16, dtype=torch.bfloat16, device=self.device)
        cb = self.scene_bias(lat_table, scene_ids, None, out=b["cbias"])
        io = L.TrainIO()
        io.b_last = self._b_last().data_ptr()
        io.training, io.dropout_off = 1, 1
        flat = samples.view(-1)
        io.gt = flat[3:].data_ptr()
        io.gt_stride = 4
        io.clamp = float(clamp)
        io.inv_n = 1.0 / float(rows_per_shape)
        io.dy_out = b["dy"].data_ptr()
        io.tile_loss = b["tile_loss"].data_ptr()
        io.bitmask = b["bitmask"].data_ptr()
        io.n_rows_pad = N
        io.tile_rows = tile_rows.data_ptr() if tile_rows is not None else None
        lat = self.lat_layer
        sums = 1 | ((1 << lat) if lat > 0 else 0)
        lib, desc, st = self.lib, self.desc, self._st()
        with torch.cuda.device(self.device):
            L.check(lib.dsdf_fused_forward(C.byref(desc), self.pack.data_ptr(), cb.data_ptr(), tile_scene.data_ptr(),
                                           flat.data_ptr(), 4, N, b["sdf"].data_ptr(), C.byref(io), st))
            L.check(lib.dsdf_fused_dgrad(C.byref(desc), self.pack.data_ptr(), C.byref(self._tb), flat.data_ptr(), 4, N, 0,
                                         0, sums, st))
            L.check(lib.dsdf_fused_latent_finish(C.byref(desc), self.pack.data_ptr(), C.byref(self._tb), tile_scene.data_ptr(), N,
                                                 lat_table.data_ptr(), scene_ids.data_ptr(), S, float(l2reg),
                                                 1.0 / float(rows_per_shape), lat_grad.data_ptr(), shape_loss.data_ptr(), st))

    # ------------------------------------------------------------------ training
    def train_buffers(self, n_rows: int, n_scenes: int) -> Dict[str, torch.Tensor]:
        """Caller-owned workspaces of the training step, sized by ``dsdf_train_buf_sizes``.  A set allocated for a larger
        (rows, scenes) is reused as it is: every size is monotonic in both, and the kernels lay the regions out from the
        row count of the call - so the chunks of ``--batch_split`` share the allocation of the largest chunk."""
        k = self._bufs_key
        if k is not None and k[0] == "train" and k[1] >= n_rows and k[2] >= n_scenes:
            return self._bufs
        key = ("train", n_rows, n_scenes)
        sizes = (C.c_size_t * 10)()
        L.check(self.lib.dsdf_train_buf_sizes(C.byref(self.desc), n_rows, n_scenes, sizes))
        names = ["acts", "dys", "bitmask", "dy", "tile_loss", "tile_part", "wg_part", "last_part", "scene_sums", "cbias"]
        self._bufs = {}
        for n, s in zip(names, sizes):
            # zero-filled once: wgrad pads layers narrower than 512 outputs to four 128-feature blocks and reads the
            # never-written columns of their dY stash
            self._bufs[n] = torch.zeros(max(int(s), 16), dtype=torch.uint8, device=self.device)
        self._bufs["sdf"] = torch.empty(n_rows, device=self.device)
        self._bufs["loss"] = torch.zeros(2, device=self.device)
        tb = TrainBufs()
        for n in names[:9]:
            setattr(tb, n, self._bufs[n].data_ptr())
        self._tb = tb
        self._bufs_key = key
        return self._bufs

    def alloc_grads(self) -> Tuple[List[torch.Tensor], L.ParamGrads]:
        """Gradient tensors in the order of ``decoder.flat_param_list()`` + the C struct pointing at them."""
        grads = [torch.zeros_like(p) for p in self.decoder.flat_param_list()]
        return grads, self.grads_struct(grads)

    def grads_struct(self, grads: List[torch.Tensor]) -> L.ParamGrads:
        PG = L.ParamGrads()
        it = iter(grads)
        for layer in range(self.spec.num_linear):
            for name, t in zip(("v", "g", "b", "ln_w", "ln_b"), self.decoder.layer_tensors(layer)):
                if t is not None:
                    getattr(PG, name)[layer] = next(it).data_ptr()
        return PG

    def loss_and_grads(self, lat_table: torch.Tensor, scene_ids: torch.Tensor, tile_scene: torch.Tensor, xyz: torch.Tensor,
                       gt: torch.Tensor, *, clamp: float, reg_lambda_eff: float, n_total: int, code_bound: Optional[float],
                       seed: int, row0: int, PG: L.ParamGrads, lat_grad: torch.Tensor, training: bool = True,
                       point_stride: int = 3, gt_stride: int = 1, n_rows: Optional[int] = None, accumulate: bool = False,
                       scene_capacity: int = 0, side_stream: Optional[torch.cuda.Stream] = None, after_scene=None,
                       step_state: Optional[torch.Tensor] = None) -> torch.Tensor:
        """train_deep_sdf.py:483-533 for one (local) batch: returns the 2-element loss tensor (total, data term); fills
        the decoder gradients behind ``PG`` and the dense latent gradient ``lat_grad`` [S_total, L] (``accumulate``: adds to
        them and to the loss instead - one chunk of ``--batch_split``).
        ``n_total`` is the GLOBAL batch size (loss normaliser, :519).  ``point_stride=4, gt_stride=4`` read the
        reference's [N,4] (xyz, sdf) sample rows in place: pass ``xyz = samples`` and ``gt = samples.view(-1)[3:]``."""
        N = int(n_rows) if n_rows is not None else xyz.shape[0]
        if N % TILE:
            raise L.DsdfError("fused training needs a multiple of 128 rows (SamplesPerScene % 128 == 0)")
        if not training:
            self._ensure_eval()
        S = scene_ids.numel()
        b = self.train_buffers(N, max(S, int(scene_capacity)))
        tb = self._tb
        cb = self.scene_bias(lat_table, scene_ids, code_bound, out=b["cbias"].view(torch.bfloat16)[:S * 2 * 512 * 16].view(S, 2, 512, 16))
        io = L.TrainIO()
        io.b_last = self._b_last().data_ptr()
        io.training = 1                              # loss / dy / gates / stash are produced either way
        io.dropout_off = int(not training)           # training=False: the same step with dropout disabled
        io.seed = seed
        io.row0 = row0
        io.step_state = step_state.data_ptr() if step_state is not None else None     # dropout keyed by the device step counter
        io.gt = gt.data_ptr()
        io.gt_stride = gt_stride
        io.clamp = float(clamp or 0.0)
        io.inv_n = 1.0 / float(n_total)
        io.dy_out = b["dy"].data_ptr()
        io.tile_loss = b["tile_loss"].data_ptr()
        io.bitmask = b["bitmask"].data_ptr()
        io.acts = b["acts"].data_ptr()
        io.n_rows_pad = N
        lib, desc, st = self.lib, self.desc, self._st()
        ev = self.kernel_events

        def mark():
            if ev is None:
                return None
            e = torch.cuda.Event(enable_timing=True)
            e.record()
            return e

        with torch.cuda.device(self.device):
            e0 = mark()
            L.check(lib.dsdf_fused_forward(C.byref(desc), self.pack.data_ptr(), cb.data_ptr(), tile_scene.data_ptr(), xyz.data_ptr(),
                                           point_stride, N, b["sdf"].data_ptr(), C.byref(io), st))
            e1 = mark()
            L.check(lib.dsdf_fused_dgrad(C.byref(desc), self.pack.data_ptr(), C.byref(tb), xyz.data_ptr(), point_stride, N,
                                         int(training), STORE_DYS_TRAIN, SUMS_ALL, st))
            e2 = mark()

            def finish(phases, stream_ptr):
                L.check(lib.dsdf_fused_finish(C.byref(desc), self.pack.data_ptr(), C.byref(self._P), C.byref(tb), tile_scene.data_ptr(), N,
                                              lat_table.data_ptr(), scene_ids.data_ptr(), S, lat_table.shape[0], float(reg_lambda_eff),
                                              1.0 / float(n_total), int(training), int(accumulate), phases, C.byref(PG),
                                              lat_grad.data_ptr(), b["loss"].data_ptr(), stream_ptr))

            if side_stream is None:
                L.check(lib.dsdf_fused_wgrad(C.byref(desc), C.byref(tb), N, st))
                e3 = mark()
                finish(3, st)
                if after_scene is not None:
                    after_scene()
            else:
                # the per-scene sums, the dense latent gradient (and whatever the caller hangs behind them: the latent Adam) need
                # only the dgrad chain's per-tile sums: they run on a second stream under the HBM-bound wgrad contraction
                main = torch.cuda.current_stream(self.device)
                side_stream.wait_stream(main)
                with torch.cuda.stream(side_stream):
                    finish(1, side_stream.cuda_stream)
                    if after_scene is not None:
                        after_scene()
                L.check(lib.dsdf_fused_wgrad(C.byref(desc), C.byref(tb), N, st))
                e3 = mark()
                main.wait_stream(side_stream)
                finish(2, st)
            e4 = mark()
        if ev is not None:
            ev.extend([("fused_forward", e0, e1), ("fused_dgrad", e1, e2), ("wgrad", e2, e3), ("finish", e3, e4)])
        return b["loss"]

    def adam(self, p: torch.Tensor, g: torch.Tensor, m: torch.Tensor, v: torch.Tensor, step: int, lr: float,
             grad_scale: float = 1.0, betas=(0.9, 0.999), eps: float = 1e-8, scale_dev: Optional[torch.Tensor] = None,
             step_state: Optional[torch.Tensor] = None) -> None:
        """torch.optim.Adam update (train_deep_sdf.py:400-411,545) of one flat fp32 tensor.  With ``step_state`` the bias
        corrections come from the device-resident step counter (``dsdf_step_advance``): replayable inside a CUDA graph."""
        if step_state is not None:
            with torch.cuda.device(self.device):
                L.check(self.lib.dsdf_adam_state(p.data_ptr(), g.data_ptr(), m.data_ptr(), v.data_ptr(), p.numel(), float(lr), betas[0],
                                                 betas[1], eps, step_state.data_ptr(), float(grad_scale),
                                                 scale_dev.data_ptr() if scale_dev is not None else None, self._st()))
            return
        with torch.cuda.device(self.device):
            L.check(self.lib.dsdf_adam(p.data_ptr(), g.data_ptr(), m.data_ptr(), v.data_ptr(), p.numel(), float(lr), betas[0], betas[1],
                                       eps, int(step), float(grad_scale), scale_dev.data_ptr() if scale_dev is not None else None,
                                       self._st()))

    def clip_coef(self, g: torch.Tensor, max_norm: float, scratch: torch.Tensor, out: torch.Tensor) -> torch.Tensor:
        """clip_grad_norm_ (train_deep_sdf.py:541-543): out[0] = ||g||, out[1] = min(1, max_norm/(||g||+1e-6))."""
        with torch.cuda.device(self.device):
            L.check(self.lib.dsdf_clip_coef(g.data_ptr(), g.numel(), float(max_norm), scratch.data_ptr(), out.data_ptr(), self._st()))
        return out
