"""CPU oracle for the DeepSDF auto-decoder hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``deepsdf_b200/`` may import this file; only
``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl
reference`` legs use it, and there only as the checker (or the CPU baseline), never as
the product path.

It is a plain fp32 PyTorch-on-CPU restatement of the reference's algorithm.  Every
function cites the reference file:line it follows (paths relative to the reference root
``mkofler96/DeepSDF``).  The one deliberate extension is an *explicit dropout mask*
argument: the reference draws its masks from torch's global RNG
(``deep_sdf/networks/deep_sdf_decoder.py:105-106``), which no other implementation can
replay, so parity runs hand the same {0,1} masks to the oracle and to the CUDA path.

Parity pin: ``oracle/make_golden.py`` imports the *real* reference ``Decoder`` (by file
path, in the build container) and the reference's own TorchScript export
``experiments/corner_spheres_only_small_network/cpp_model.pt`` and writes
``tests/golden/*.npz``; ``tests/test_oracle_golden.py`` checks this restatement against
those vectors (max abs diff <= 5e-6, fp32 summation-order noise).  ``reconstruct`` (latent-only optimisation) is the
exception: the script is absent from the reference fork, so that function restates the
upstream facebookresearch/DeepSDF ``reconstruct.py`` defaults and is **parity unpinned**.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence

import numpy as np
import torch
import torch.nn.functional as F


def kernel_dropout_scale(p: float) -> float:
    """1 / (1 - effective p) of the CUDA kernels' 8-bit Philox dropout (csrc/common.cuh dropout_scale): the keep
    probability is (256 - round(256 p)) / 256.  Not reference semantics - only for tests that share the kernels' masks."""
    thr = min(255, max(0, int(p * 256.0 + 0.5)))
    return 256.0 / (256 - thr)


# --------------------------------------------------------------------------------------
# network description (deep_sdf_decoder.py:10-73)
# --------------------------------------------------------------------------------------
@dataclass
class OracleSpec:
    latent_size: int
    dims: Sequence[int]
    geom_dimension: int = 3
    dropout: Optional[Sequence[int]] = None
    dropout_prob: float = 0.0
    norm_layers: Sequence[int] = ()
    latent_in: Sequence[int] = ()
    weight_norm: bool = False
    xyz_in_all: Optional[bool] = None
    use_tanh: bool = False
    latent_dropout: bool = False
    # Inverted-dropout scale of the survivors.  None = the reference's F.dropout: 1 / (1 - p) (deep_sdf_decoder.py:105-106).
    # The CUDA kernels draw 8-bit Philox numbers, so their keep probability is (256 - round(256 p)) / 256 and they scale
    # by its reciprocal; a test that hands the kernels' masks to the oracle sets this to the same number
    # (kernel_dropout_scale) so that both sides evaluate the same function.
    dropout_scale: Optional[float] = None
    # Only read by decoder_forward_bf16_operands (the emulation of the CUDA chain's operand rounding): True = the dropout
    # scale of hidden layer l is multiplied into the activation columns of layer l+1's weights BEFORE their bf16 rounding
    # and the activations are rounded unscaled - the arithmetic of the round-2 kernels (csrc/fused_host.cu, training pack).
    fold_scale_into_weights: bool = False
    # derived
    full_dims: List[int] = field(default_factory=list)
    in_dims: List[int] = field(default_factory=list)
    out_dims: List[int] = field(default_factory=list)

    def __post_init__(self):
        # deep_sdf_decoder.py:29  dims = [latent_size + geom_dimension] + dims + [1]
        d = [self.latent_size + self.geom_dimension] + list(self.dims) + [1]
        self.full_dims = d
        self.in_dims, self.out_dims = [], []
        n = len(d)
        for layer in range(n - 1):
            # deep_sdf_decoder.py:42-48
            if layer + 1 in self.latent_in:
                out_dim = d[layer + 1] - d[0]
            else:
                out_dim = d[layer + 1]
                if self.xyz_in_all and layer != n - 2:
                    out_dim -= self.geom_dimension
            self.in_dims.append(d[layer])
            self.out_dims.append(out_dim)

    @property
    def num_linear(self) -> int:
        return len(self.full_dims) - 1

    def drop_scale(self) -> float:
        return float(self.dropout_scale) if self.dropout_scale is not None else 1.0 / (1.0 - self.dropout_prob)

    def latent_drop_scale(self) -> float:
        # latent_dropout uses p = 0.2 whatever dropout_prob says (deep_sdf_decoder.py:81); with the kernels' masks the
        # scale follows the same 8-bit quantisation as dropout_scale
        return kernel_dropout_scale(0.2) if self.dropout_scale is not None else 1.0 / (1.0 - 0.2)

    def is_weight_normed(self, layer: int) -> bool:
        # deep_sdf_decoder.py:50
        return bool(self.weight_norm) and layer in self.norm_layers

    def has_layernorm(self, layer: int) -> bool:
        # deep_sdf_decoder.py:59-64 (constructed for every such layer, applied only to hidden ones :97-103)
        return (not self.weight_norm) and self.norm_layers is not None and layer in self.norm_layers

    def param_names(self) -> List[str]:
        """state_dict key order of the reference module (what ``named_parameters`` yields)."""
        names = []
        for layer in range(self.num_linear):
            if self.is_weight_normed(layer):
                names += [f"lin{layer}.bias",
                          f"lin{layer}.parametrizations.weight.original0",
                          f"lin{layer}.parametrizations.weight.original1"]
            else:
                names += [f"lin{layer}.weight", f"lin{layer}.bias"]
            if self.has_layernorm(layer):
                names += [f"bn{layer}.weight", f"bn{layer}.bias"]
        return names

    def param_shapes(self) -> Dict[str, tuple]:
        shapes = {}
        for layer in range(self.num_linear):
            o, i = self.out_dims[layer], self.in_dims[layer]
            if self.is_weight_normed(layer):
                shapes[f"lin{layer}.bias"] = (o,)
                shapes[f"lin{layer}.parametrizations.weight.original0"] = (o, 1)
                shapes[f"lin{layer}.parametrizations.weight.original1"] = (o, i)
            else:
                shapes[f"lin{layer}.weight"] = (o, i)
                shapes[f"lin{layer}.bias"] = (o,)
            if self.has_layernorm(layer):
                shapes[f"bn{layer}.weight"] = (o,)
                shapes[f"bn{layer}.bias"] = (o,)
        return shapes


def make_params(spec: OracleSpec, seed: int, scale: float = 2.4) -> Dict[str, torch.Tensor]:
    """Deterministic, torch-version-independent parameters (numpy RandomState).  Same
    distribution family as ``nn.Linear``'s default init (uniform, bound ~ 1/sqrt(in)) but
    with the weight bound widened by ``scale`` (~He gain) so that the signal does not die
    out over 8 layers and outputs span a useful part of (-1, 1), like a trained network;
    weight-norm g starts from the row norm of v (as ``weight_norm`` does) and is perturbed."""
    rng = np.random.RandomState(seed)
    p: Dict[str, torch.Tensor] = {}
    for layer in range(spec.num_linear):
        o, i = spec.out_dims[layer], spec.in_dims[layer]
        bound = scale / math.sqrt(i)
        w = rng.uniform(-bound, bound, size=(o, i)).astype(np.float32)
        b = rng.uniform(-bound / scale, bound / scale, size=(o,)).astype(np.float32)
        if spec.is_weight_normed(layer):
            g = np.linalg.norm(w.astype(np.float64), axis=1, keepdims=True).astype(np.float32)
            # perturb g so that g != ||v|| (a trained network's state)
            g = g * rng.uniform(0.8, 1.25, size=g.shape).astype(np.float32)
            p[f"lin{layer}.bias"] = torch.from_numpy(b)
            p[f"lin{layer}.parametrizations.weight.original0"] = torch.from_numpy(g)
            p[f"lin{layer}.parametrizations.weight.original1"] = torch.from_numpy(w)
        else:
            p[f"lin{layer}.weight"] = torch.from_numpy(w)
            p[f"lin{layer}.bias"] = torch.from_numpy(b)
        if spec.has_layernorm(layer):
            p[f"bn{layer}.weight"] = torch.from_numpy(rng.uniform(0.5, 1.5, size=(o,)).astype(np.float32))
            p[f"bn{layer}.bias"] = torch.from_numpy(rng.uniform(-0.1, 0.1, size=(o,)).astype(np.float32))
    return p


# --------------------------------------------------------------------------------------
# forward (deep_sdf_decoder.py:76-111)
# --------------------------------------------------------------------------------------
def effective_weight(spec: OracleSpec, params: Dict[str, torch.Tensor], layer: int) -> torch.Tensor:
    """``torch._weight_norm(v, g, dim=0)``: W = g * v / ||v||_2 taken per output row
    (deep_sdf_decoder.py:50-55; the parametrisation is re-evaluated on every call)."""
    if spec.is_weight_normed(layer):
        g = params[f"lin{layer}.parametrizations.weight.original0"]
        v = params[f"lin{layer}.parametrizations.weight.original1"]
        return v * (g / v.norm(dim=1, keepdim=True))
    return params[f"lin{layer}.weight"]


def decoder_forward(spec: OracleSpec, params: Dict[str, torch.Tensor], inp: torch.Tensor,
                    training: bool = False,
                    masks: Optional[Dict[int, torch.Tensor]] = None,
                    latent_mask: Optional[torch.Tensor] = None) -> torch.Tensor:
    """deep_sdf_decoder.py:76-111.  ``masks[layer]`` is the {0,1} keep-mask of hidden
    layer ``layer`` ([N, out_dim]); ``latent_mask`` the keep-mask of the latent columns
    (latent_dropout, p fixed to 0.2 at :81).  Masks are only consulted when ``training``."""
    D = spec.geom_dimension
    xyz = inp[:, -D:]                                                     # :77
    if inp.shape[1] > D and spec.latent_dropout:                          # :79
        lat = inp[:, :-D]
        if training:
            if latent_mask is None:
                raise ValueError("training with latent_dropout needs latent_mask")
            lat = lat * latent_mask * spec.latent_drop_scale()             # :81 F.dropout(p=0.2)
        x = torch.cat([lat, xyz], 1)                                      # :82
    else:
        x = inp                                                           # :84
    n = spec.num_linear
    for layer in range(n):                                                # :86
        if layer in spec.latent_in:                                       # :88
            x = torch.cat([x, inp], 1)
        elif layer != 0 and spec.xyz_in_all:                              # :90
            x = torch.cat([x, xyz], 1)
        W = effective_weight(spec, params, layer)
        x = F.linear(x, W, params[f"lin{layer}.bias"])                    # :92
        if layer == n - 1 and spec.use_tanh:                              # :94
            x = torch.tanh(x)
        if layer < n - 1:                                                 # :96
            if spec.has_layernorm(layer):                                 # :97-103
                x = F.layer_norm(x, (x.shape[1],), params[f"bn{layer}.weight"], params[f"bn{layer}.bias"])
            x = torch.relu(x)                                             # :104
            if spec.dropout is not None and layer in spec.dropout and training and spec.dropout_prob > 0:
                if masks is None or layer not in masks:                   # :105-106
                    raise ValueError(f"training forward needs an explicit mask for layer {layer}")
                x = x * masks[layer] * spec.drop_scale()
    return torch.tanh(x)                                                  # :108-109 (self.th always exists)


def decode_sdf(spec, params, latent_vector, queries, **kw):
    """deep_sdf/utils.py:54-65."""
    if latent_vector is None:
        inputs = queries
    else:
        inputs = torch.cat([latent_vector.reshape(1, -1).expand(queries.shape[0], -1), queries], 1)
    return decoder_forward(spec, params, inputs, **kw)


def input_gradient(spec, params, inp: torch.Tensor) -> torch.Tensor:
    """G[N, L+D] = dSDF/d[latent, xyz] in eval mode: what ``torch.autograd.functional.jvp``
    through the decoder contracts with its tangents (deep_sdf/mesh.py:405-422); rows are
    independent so one reverse pass with a ones cotangent gives every row's gradient."""
    x = inp.detach().clone().requires_grad_(True)
    y = decoder_forward(spec, params, x, training=False)
    (g,) = torch.autograd.grad(y.sum(), x)
    return g


# --------------------------------------------------------------------------------------
# training step (train_deep_sdf.py:481-545)
# --------------------------------------------------------------------------------------
def embedding_renorm_(table: torch.Tensor, rows: torch.Tensor, max_norm: Optional[float]) -> None:
    """``nn.Embedding(max_norm=CodeBound)`` lookup side effect (train_deep_sdf.py:385,509):
    every *touched* row with ||z||_2 > max_norm is rescaled in place, without grad, by
    max_norm / (||z|| + 1e-7) (ATen ``embedding_renorm_``)."""
    if max_norm is None:
        return
    with torch.no_grad():
        for r in torch.unique(rows).tolist():
            nrm = float(table[r].norm(2))
            if nrm > max_norm:
                table[r] *= max_norm / (nrm + 1e-7)


def train_loss_and_grads(spec: OracleSpec, params: Dict[str, torch.Tensor], lat_table: torch.Tensor,
                         scene_of_row: torch.Tensor, xyz: torch.Tensor, sdf_gt: torch.Tensor,
                         clamp_dist: Optional[float], code_reg_lambda: Optional[float], epoch: int,
                         masks: Optional[Dict[int, torch.Tensor]] = None, training: bool = True,
                         code_bound: Optional[float] = None, n_total: Optional[int] = None):
    """One optimiser-step's loss and gradients (batch_split = 1), train_deep_sdf.py:483-533.
    ``n_total`` overrides the loss normaliser (the whole batch's row count, :485,519) when the rows given are one
    shard of a larger batch - the same role it plays under ``batch_split``.

    Returns (loss, {param name: grad}, lat_grad[S_total, L] dense, pred_sdf_raw[N,1]).
    ``lat_table`` is modified in place by the max_norm renorm exactly like the reference.
    """
    embedding_renorm_(lat_table, scene_of_row, code_bound)                  # :509 (lookup side effect)
    P = {k: v.detach().clone().requires_grad_(True) for k, v in params.items()}
    Z = lat_table.detach().clone().requires_grad_(True)
    n_total = xyz.shape[0] if n_total is None else n_total                # :485
    gt = sdf_gt.reshape(-1, 1)
    if clamp_dist is not None:                                            # :492-493
        gt = torch.clamp(gt, -clamp_dist, clamp_dist)
    batch_vecs = Z[scene_of_row]                                          # :509
    inp = torch.cat([batch_vecs, xyz], dim=1)                             # :511
    pred = decoder_forward(spec, P, inp, training=training, masks=masks)  # :514
    pred_c = torch.clamp(pred, -clamp_dist, clamp_dist) if clamp_dist is not None else pred  # :516-517
    loss = (pred_c - gt).abs().sum() / n_total                            # :398,519 (L1 sum / N)
    if code_reg_lambda is not None:                                       # :523-531
        l2 = torch.sum(torch.norm(batch_vecs, dim=1))
        loss = loss + (code_reg_lambda * min(1, epoch / 100) * l2) / n_total
    loss.backward()                                                       # :533
    grads = {k: v.grad.detach() for k, v in P.items()}
    return float(loss.detach()), grads, Z.grad.detach(), pred.detach()


def adam_step_(param: torch.Tensor, grad: torch.Tensor, m: torch.Tensor, v: torch.Tensor, step: int,
               lr: float, beta1: float = 0.9, beta2: float = 0.999, eps: float = 1e-8) -> None:
    """``torch.optim.Adam`` single-tensor update with default hyper-parameters
    (train_deep_sdf.py:400-411,545): dense, no weight decay, no amsgrad."""
    m.mul_(beta1).add_(grad, alpha=1 - beta1)
    v.mul_(beta2).addcmul_(grad, grad, value=1 - beta2)
    bc1 = 1 - beta1 ** step
    bc2 = 1 - beta2 ** step
    denom = (v.sqrt() / math.sqrt(bc2)).add_(eps)
    param.addcdiv_(m, denom, value=-lr / bc1)


def step_lr(initial: float, interval: int, factor: float, epoch: int) -> float:
    """train_deep_sdf.py:42-44 StepLearningRateSchedule."""
    return initial * (factor ** (epoch // interval))


# --------------------------------------------------------------------------------------
# grid evaluation (deep_sdf/mesh.py:37-73)
# --------------------------------------------------------------------------------------
def grid_points(N: int) -> torch.Tensor:
    """deep_sdf/mesh.py:37-53: linear index = (i*N + j)*N + k ; coord = idx*2/(N-1) - 1,
    column 0 <- i (slowest), column 2 <- k (fastest)."""
    voxel_origin = -1.0
    voxel_size = 2.0 / (N - 1)
    idx = torch.arange(0, N ** 3, 1, dtype=torch.long)
    s = torch.zeros(N ** 3, 3)
    s[:, 2] = idx % N
    s[:, 1] = (idx.long() // N) % N          # mesh.py:46 (float division + .long() in the reference)
    s[:, 0] = ((idx.long() // N) // N) % N   # mesh.py:47
    return s * voxel_size + voxel_origin     # mesh.py:50-52


def create_mesh_sdf(spec, params, latent_vec: torch.Tensor, N: int, max_batch: int = 32 ** 3) -> torch.Tensor:
    """The chunk loop of ``create_mesh`` (deep_sdf/mesh.py:59-73) without marching cubes."""
    pts = grid_points(N)
    out = torch.zeros(N ** 3)
    head = 0
    with torch.no_grad():
        while head < N ** 3:
            q = pts[head:head + max_batch]
            out[head:head + max_batch] = decode_sdf(spec, params, latent_vec, q).squeeze(1)
            head += max_batch
    return out.reshape(N, N, N)


# --------------------------------------------------------------------------------------
# latent-only reconstruction (upstream reconstruct.py; absent from the fork -> parity unpinned)
# --------------------------------------------------------------------------------------
def reconstruct(spec, params, xyz: torch.Tensor, sdf_gt: torch.Tensor, num_iterations: int = 800,
                num_samples: int = 8000, clamp_dist: float = 0.1, lr: float = 5e-3, l2reg: bool = True,
                init_std: float = 0.01, seed: int = 0, sample_idx: Optional[torch.Tensor] = None):
    """PARITY UNPINNED.  Restates upstream facebookresearch/DeepSDF ``reconstruct.py``
    (the fork only keeps README.md:134-142 and the helpers deep_sdf/data.py:66-71,113-139):
    latent ~ N(0, init_std) [1,L]; Adam(lr); lr *= 0.1 every num_iterations/2; per iteration a
    subsample of ``num_samples`` rows, gt and pred clamped to +-clamp_dist, L1 *mean* loss
    + 1e-4 * mean(z^2); decoder in eval mode and frozen.
    ``sample_idx[it]`` (optional, [iters, num_samples] long) fixes the subsamples."""
    g = torch.Generator().manual_seed(seed)
    z = (torch.randn(1, spec.latent_size, generator=g) * init_std).requires_grad_(True)
    opt = torch.optim.Adam([z], lr=lr)
    decreased_by, adjust_lr_every = 10, int(num_iterations / 2)
    losses = []
    for e in range(num_iterations):
        cur_lr = lr * ((1 / decreased_by) ** (e // adjust_lr_every))
        for pg in opt.param_groups:
            pg["lr"] = cur_lr
        if sample_idx is not None:
            idx = sample_idx[e]
        else:
            idx = torch.randint(0, xyz.shape[0], (num_samples,), generator=g)
        q = xyz[idx]
        gt = torch.clamp(sdf_gt[idx].reshape(-1, 1), -clamp_dist, clamp_dist)
        opt.zero_grad()
        inp = torch.cat([z.expand(q.shape[0], -1), q], 1)
        pred = torch.clamp(decoder_forward(spec, params, inp, training=False), -clamp_dist, clamp_dist)
        loss = (pred - gt).abs().mean()
        if l2reg:
            loss = loss + 1e-4 * torch.mean(z.pow(2))
        loss.backward()
        opt.step()
        losses.append(float(loss.detach()))
    return z.detach(), losses


# --------------------------------------------------------------------------------------
# synthetic data: analytic SDFs (sdf_sampler/microstructures.py:3-36)
# --------------------------------------------------------------------------------------
def corner_spheres_sdf(xyz: np.ndarray, radius: float, limit: float = 1.0) -> np.ndarray:
    """sdf_sampler/microstructures.py:23-36 CornerSpheresSDF.SDF."""
    out = np.linalg.norm(xyz, axis=1, ord=np.inf) - limit
    corners = np.array(np.meshgrid([-1, 1], [-1, 1], [-1, 1])).T.reshape(-1, 3)
    for c in corners:
        sphere_like = np.linalg.norm(xyz - c, axis=1, ord=3) - radius
        out = np.maximum(out, -sphere_like)
    return out.reshape(-1, 1)


def cross_ms_sdf(xyz: np.ndarray, radius: float) -> np.ndarray:
    """sdf_sampler/microstructures.py:3-20 CrossMsSDF.SDF."""
    out = np.linalg.norm(xyz, axis=1, ord=np.inf)
    for a, b in ((1, 2), (0, 2), (0, 1)):
        out = np.minimum(out, np.sqrt(xyz[:, a] ** 2 + xyz[:, b] ** 2) - radius)
    return out.reshape(-1, 1)


# --------------------------------------------------------------------------------------
# test-input conditioning (not part of the reference)
# --------------------------------------------------------------------------------------
def well_conditioned_batch(spec: OracleSpec, params, lat_table: torch.Tensor, scene_of_row: torch.Tensor,
                           masks: Optional[Dict[int, torch.Tensor]], clamp_dist: float, seed: int,
                           margin: float = 6e-3, candidates: int = 4):
    """Synthetic (xyz, sdf_gt) for gradient-parity tests of a reduced-precision implementation.

    The clamped-L1 loss has a discontinuous derivative: dL/dpred = sign(pred - gt) inside the clamp band, 0 outside
    (train_deep_sdf.py:516-519).  A bf16 forward that is 1e-3 off can flip that sign, or the in-band decision, for rows
    whose prediction happens to sit within 1e-3 of gt or of +-clamp, and each flipped row changes the gradient by O(1)
    of its contribution - noise that says nothing about the backward kernels.  This helper picks, per row position (the
    dropout masks are tied to the position), an xyz among ``candidates`` draws whose fp32 prediction is at least
    ``margin`` away from +-clamp, and a gt at least ``margin`` (and at most 0.05) away from the prediction.
    ``lat_table`` must already be renormalised (embedding_renorm_)."""
    rng = np.random.RandomState(seed)
    n = scene_of_row.shape[0]
    D = spec.geom_dimension
    xyz = torch.zeros(n, D)
    pred = torch.zeros(n)
    done = torch.zeros(n, dtype=torch.bool)
    with torch.no_grad():
        for _ in range(candidates):
            cand = torch.from_numpy(rng.uniform(-1, 1, (n, D)).astype(np.float32))
            p = decoder_forward(spec, params, torch.cat([lat_table[scene_of_row], cand], 1), training=masks is not None,
                                masks=masks).squeeze(1)
            ok = ((p.abs() - clamp_dist).abs() >= margin) & ~done
            xyz[ok] = cand[ok]
            pred[ok] = p[ok]
            done |= ok
        # positions never satisfied keep their last candidate (rare; they are few enough not to matter)
        xyz[~done] = cand[~done]
        pred[~done] = p[~done]
    sign = torch.from_numpy(rng.choice([-1.0, 1.0], n).astype(np.float32))
    gap = torch.from_numpy(rng.uniform(margin, 0.05, n).astype(np.float32))
    pc = pred.clamp(-clamp_dist, clamp_dist)
    gt = pc + sign * gap
    over = gt.abs() > clamp_dist          # gt is clamped too (train_deep_sdf.py:493): keep it inside the band
    gt[over] = (pc - sign * gap)[over]
    return xyz, gt, int((~done).sum())


# --------------------------------------------------------------------------------------
# the same step at the CUDA path's operand precision (checker for the backward kernels)
# --------------------------------------------------------------------------------------
class _RoundBF16(torch.autograd.Function):
    """Round to bf16 in the forward, identity in the backward (straight-through)."""

    @staticmethod
    def forward(ctx, x):
        return x.to(torch.bfloat16).float()

    @staticmethod
    def backward(ctx, g):
        return g


def decoder_forward_bf16_operands(spec: OracleSpec, params, lat_rows: torch.Tensor, xyz: torch.Tensor,
                                  training: bool = False, masks: Optional[Dict[int, torch.Tensor]] = None,
                                  gates: Optional[Dict[int, torch.Tensor]] = None) -> torch.Tensor:
    """``decoder_forward`` (deep_sdf_decoder.py:76-111) with the operand rounding of the fused tcgen05 path: hidden
    activations and the hidden-column weight blocks are rounded to bf16 before each product (fp32 accumulate); the latent
    and xyz columns of lin0 / the latent_in layer, all biases, and the last layer stay fp32.  ``lat_rows`` = the gathered
    latent rows [N, L].  Differentiable (straight-through rounding), so autograd through it gives the EXACT gradient of
    the function the kernels evaluate.  Why it exists: ReLU gates whose fp32 pre-activation is within bf16 noise of zero
    open or close differently under bf16 operands; each flipped gate changes its gradient entry by 100 %, so a gradient
    comparison against the pure fp32 forward measures the gate flips (about 3-4 % relative norm per layer with random
    weights, independent of the batch size), not the backward kernels.  Even two bf16-operand forwards that differ only
    in fp32 summation order disagree on a few gates (0.5-0.8 % relative norm per gradient, measured), so for the 1e-2
    gradient check the caller can hand over the gates the CUDA forward actually took, like the shared dropout masks:
    ``gates[l]`` = {0,1} [N, out_dim] = relu gate AND dropout keep of hidden layer l; then h = pre * gate * scale.
    Restricted to the fused path's network class (weight-norm or plain layers, one latent_in layer, no LayerNorm /
    xyz_in_all / latent_dropout)."""
    L, D = spec.latent_size, spec.geom_dimension
    nh = spec.num_linear - 1
    a = None
    h = None
    fold = bool(spec.fold_scale_into_weights)
    pend = 1.0               # fold: scale of the previous layer's dropout, owed to this layer's activation columns
    for l in range(nh):
        W = effective_weight(spec, params, l)
        b = params[f"lin{l}.bias"]
        if l == 0:
            h = lat_rows @ W[:, :L].t() + b + xyz @ W[:, L:L + D].t()
        elif l in spec.latent_in:
            xw = spec.in_dims[l] - (L + D)
            h = a[:, :xw] @ _RoundBF16.apply(W[:, :xw] * pend).t() + (lat_rows @ W[:, xw:xw + L].t() + b) + xyz @ W[:, xw + L:].t()
        else:
            h = a @ _RoundBF16.apply(W * pend).t() + b
        drop = training and spec.dropout is not None and l in spec.dropout and spec.dropout_prob > 0
        sc = spec.drop_scale() if drop else 1.0
        if gates is not None:
            h = h * gates[l]
        else:
            h = torch.relu(h)
            if drop:
                h = h * masks[l]
        if fold:
            pend = sc
        else:
            h = h * sc
        a = _RoundBF16.apply(h)
    y = h @ (effective_weight(spec, params, nh) * (pend if fold else 1.0)).t() + params[f"lin{nh}.bias"]
    if spec.use_tanh:
        y = torch.tanh(y)
    return torch.tanh(y)


def train_loss_and_grads_bf16_operands(spec: OracleSpec, params, lat_table, scene_of_row, xyz, sdf_gt, clamp_dist,
                                       reg_lambda_eff: float, n_total: int, masks=None, training: bool = True, gates=None):
    """train_deep_sdf.py:483-533 evaluated through ``decoder_forward_bf16_operands``.  ``lat_table`` must already be
    renormalised; ``reg_lambda_eff`` = CodeRegularizationLambda * min(1, epoch/100) (or 0); ``n_total`` the global rows."""
    P = {k: v.detach().clone().requires_grad_(True) for k, v in params.items()}
    Z = lat_table.detach().clone().requires_grad_(True)
    bv = Z[scene_of_row]
    pred = decoder_forward_bf16_operands(spec, P, bv, xyz, training=training, masks=masks, gates=gates)
    gt = sdf_gt.reshape(-1, 1)
    if clamp_dist is not None:
        gt = gt.clamp(-clamp_dist, clamp_dist)
        pred_c = pred.clamp(-clamp_dist, clamp_dist)
    else:
        pred_c = pred
    loss = (pred_c - gt).abs().sum() / n_total
    if reg_lambda_eff:
        loss = loss + reg_lambda_eff * torch.sum(torch.norm(bv, dim=1)) / n_total
    loss.backward()
    return float(loss.detach()), {k: v.grad.detach() for k, v in P.items()}, Z.grad.detach(), pred.detach()
