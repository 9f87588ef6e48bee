"""Drop-in ``Decoder`` for ``deep_sdf.networks.deep_sdf_decoder.Decoder``.

Same constructor keywords, same parameter / ``state_dict`` names (``lin{l}.bias``,
``lin{l}.parametrizations.weight.original0|1`` or ``lin{l}.weight``, ``bn{l}.*``), same
attributes (``geom_dimension``, ``num_layers``, ``latent_in`` ...), same ``forward(input[N, L+D])
-> [N, 1]`` contract, honours ``train()/eval()``, differentiable w.r.t. its input and its
parameters, survives ``torch.nn.DataParallel`` on one device.  Reference:
deep_sdf/networks/deep_sdf_decoder.py:9-111.

The compute is NOT PyTorch: ``forward``/``backward`` call the C-ABI library
(``include/dsdf.h``) - hand-written sm_100a kernels.  There is no CPU path and no fallback.

Additional (fused, scene-mode) entry points used by the build's own trainer / grid evaluator:
``decode_scenes`` (see ``deepsdf_b200.engine``).
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, List, Optional

import torch
import torch.nn as nn

from .. import _lib as L
from ..spec import NetSpec

_WS_SINGLE_PASS_LIMIT = 6 << 30     # keep one workspace alive for backward below this size
_CHUNK_ROWS = 1 << 17


class Decoder(nn.Module):
    def __init__(
        self,
        latent_size,
        dims,
        geom_dimension,
        dropout=None,
        dropout_prob=0.0,
        norm_layers=(),
        latent_in=(),
        weight_norm=False,
        xyz_in_all=None,
        use_tanh=False,
        latent_dropout=False,
    ):
        super().__init__()
        self.spec = NetSpec(latent_size, list(dims), geom_dimension, dropout, dropout_prob, norm_layers, latent_in,
                            weight_norm, xyz_in_all, use_tanh, latent_dropout)
        s = self.spec
        # public attributes the reference sets (deep_sdf_decoder.py:31-40, 66-72)
        self.num_layers = s.num_layers
        self.geom_dimension = geom_dimension
        self.norm_layers = norm_layers
        self.latent_in = latent_in
        self.latent_dropout = latent_dropout
        self.xyz_in_all = xyz_in_all
        self.weight_norm = weight_norm
        self.use_tanh = use_tanh
        self.dropout_prob = dropout_prob
        self.dropout = dropout
        self.latent_size = latent_size
        # parameter containers with the reference's names; they only hold tensors, their forward() is never used
        for layer in range(s.num_linear):
            lin = nn.Linear(s.in_dims[layer], s.out_dims[layer])
            if s.is_weight_normed(layer):
                lin = nn.utils.parametrizations.weight_norm(lin)
            setattr(self, "lin" + str(layer), lin)
            if s.has_layernorm(layer):
                setattr(self, "bn" + str(layer), nn.LayerNorm(s.out_dims[layer]))
        self._desc = s.desc()

    # ------------------------------------------------------------------ parameter plumbing
    def layer_tensors(self, layer: int):
        """(v_or_weight, g_or_None, bias, ln_w, ln_b) of one layer, as the live Parameters."""
        lin = getattr(self, "lin" + str(layer))
        if self.spec.is_weight_normed(layer):
            g = lin.parametrizations.weight.original0
            v = lin.parametrizations.weight.original1
        else:
            g, v = None, lin.weight
        ln_w = ln_b = None
        if self.spec.has_layernorm(layer):
            bn = getattr(self, "bn" + str(layer))
            ln_w, ln_b = bn.weight, bn.bias
        return v, g, lin.bias, ln_w, ln_b

    def flat_param_list(self) -> List[torch.Tensor]:
        out = []
        for layer in range(self.spec.num_linear):
            out.extend(t for t in self.layer_tensors(layer) if t is not None)
        return out

    def flat_param_names(self) -> List[str]:
        """state_dict names in the order of ``flat_param_list()``."""
        ids = {id(p): n for n, p in self.named_parameters()}
        return [ids[id(t)] for t in self.flat_param_list()]

    def c_params(self, tensors: Optional[List[torch.Tensor]] = None) -> L.Params:
        """Fill the C struct from the live parameters (or from a same-order tensor list)."""
        P = L.Params()
        it = iter(tensors) if tensors is not None else None
        for layer in range(self.spec.num_linear):
            live = self.layer_tensors(layer)
            vals = []
            for t in live:
                if t is None:
                    vals.append(None)
                else:
                    vals.append(next(it) if it is not None else t)
            v, g, b, ln_w, ln_b = vals
            for name, t in (("v", v), ("g", g), ("b", b), ("ln_w", ln_w), ("ln_b", ln_b)):
                if t is not None:
                    if t.dtype != torch.float32 or not t.is_contiguous():
                        raise L.DsdfError("decoder parameters must be contiguous fp32")
                    getattr(P, name)[layer] = t.data_ptr()
        return P

    # ------------------------------------------------------------------ forward
    fused_eval = True      # eval-mode forward of fused-class networks with CodeLength <= 16 on the tensor-core chain
    fused_exact = False    # consumers with one code per call (decode_sdf, create_mesh): True selects the split-fp16 high-precision
                           # tensor-core forward (<= 1e-5 vs fp32, ~4x the time of the bf16 chain; csrc/chain_exact.cu)

    def _fused_rows_supported(self) -> bool:
        ok = getattr(self, "_fused_rows_ok", None)
        if ok is None:
            ok = L.lib().dsdf_fused_rows_supported(C.byref(self._desc)) == 1
            object.__setattr__(self, "_fused_rows_ok", ok)
        return ok

    def _small_supported(self) -> bool:
        ok = getattr(self, "_small_ok", None)
        if ok is None:
            ok = L.lib().dsdf_small_supported(C.byref(self._desc)) == 1
            object.__setattr__(self, "_small_ok", ok)
        return ok

    def forward(self, input):
        if input.dim() != 2 or input.shape[1] != self.spec.full_dims[0]:
            raise RuntimeError(
                f"Decoder expects input [N, {self.spec.full_dims[0]}] (latent ++ xyz), got {tuple(input.shape)}")
        L.require_cuda(input, *self.flat_param_list())
        if torch.jit.is_tracing() and not self.training:
            # torch.jit.trace (create_libtorch_executable.py:19): one registered op instead of an autograd.Function over ctypes
            from ..torch_op import decoder_eval, encode_spec
            return decoder_eval(input, list(self.flat_param_list()), encode_spec(self.spec), float(self.spec.dropout_prob))
        if not self.training and self.fused_eval and input.shape[0] > 0 and self._fused_rows_supported():
            # eval-mode boundary call of an 8x512-class network with a short per-row code (CodeLength <= 16): the bf16
            # tcgen05 chain in row-latent mode instead of the fp32 CUDA-core path (set ``decoder.fused_eval = False`` for
            # the exact fp32 mode)
            y = _FusedRowsFn.apply(self, input)
            params = self.flat_param_list()
            if torch.is_grad_enabled() and any(p.requires_grad for p in params):
                # parameter gradients (rare in eval mode) hang on a separate, zero-valued branch of the graph: autograd walks it
                # only when a parameter's gradient is actually asked for, so autograd.grad(sdf, xyz) / the jvp loop of
                # deep_sdf/mesh.py:405-422 on a loaded model (whose parameters still require grad) never pays for them
                y = y + _ParamBranchFn.apply(self, input.detach(), *params)
            return y
        if not self.training and self.fused_eval and input.shape[0] > 0 and self._small_supported():
            # eval-mode boundary call of a small network (hidden widths <= 128: the 4x32 / 4x64 / 6x128 decoders of the reference's
            # optimisation runs): ONE fp32 launch for the SDF - and for dSDF/dinput when the input requires grad - instead of
            # the per-layer generic kernels; same fp32 arithmetic up to summation order (~1e-6)
            y = _SmallRowsFn.apply(self, input)
            params = self.flat_param_list()
            if torch.is_grad_enabled() and any(p.requires_grad for p in params):
                y = y + _ParamBranchFn.apply(self, input.detach(), *params)
            return y
        seed = 0
        if self.training and (self.spec.uses_dropout() or self.spec.latent_dropout):
            # one Philox key per call, drawn from torch's CPU generator -> torch.manual_seed makes runs repeatable.  Only
            # when the spec has dropout: a network without it must not disturb the caller's RNG stream (the reference's
            # DataLoader shuffle draws from the same generator)
            seed = int(torch.randint(0, 2 ** 62, (1,)).item())
        params = self.flat_param_list()
        return _DecoderFn.apply(self, input, bool(self.training), seed, *params)


class _SecondOrderGuard(torch.autograd.Function):
    """Identity on the per-row input gradient G, tied to the decoder's input and parameters in the autograd graph: G is a
    kernel output, its own derivatives (second order in x or in the parameters) are not provided - asking for them must be
    an error, not a silent zero."""

    @staticmethod
    def forward(ctx, G, *deps):
        return G.view_as(G)

    @staticmethod
    def backward(ctx, g):
        raise RuntimeError("deepsdf_b200.Decoder: second-order derivatives (d/dx or d/dparams of dSDF/dinput) are not provided; "
                           "only the first-order double-backward trick of torch.autograd.functional.jvp (deep_sdf/mesh.py:405-422) "
                           "and first-order parameter gradients under create_graph=True are")


class _FusedRowsFn(torch.autograd.Function):
    """Eval-mode ``decoder(input[N, L+3])`` of an 8x512-class network (CodeLength <= 16, or a long code such as 256) on the
    fused tcgen05 chain in row-latent mode (bf16 operands, fp32 accumulate: SDF within 2e-3 of the fp32 arithmetic).  The
    input gradient comes from the fused dgrad chain + input-gradient GEMM (one ``G`` per call, also under the
    double-backward trick of ``torch.autograd.functional.jvp``).  Parameter gradients live on ``_ParamBranchFn``."""

    @staticmethod
    def forward(ctx, module, inp):
        from ..utils import fused_engine_for
        eng = fused_engine_for(module)
        x = inp.detach().to(torch.float32).contiguous()
        sdf, _ = eng.sdf_and_input_grad_rows(x, with_grad=False)
        ctx.module = module
        ctx.save_for_backward(x)
        ctx.inp_ref = inp if inp.requires_grad else None
        return sdf.unsqueeze(1)

    @staticmethod
    def backward(ctx, grad_out):
        from ..utils import fused_engine_for
        (x,) = ctx.saved_tensors
        module = ctx.module
        if not ctx.needs_input_grad[1]:
            return None, None
        N = x.shape[0]
        deps = ([ctx.inp_ref] if ctx.inp_ref is not None else []) + [p for p in module.flat_param_list() if p.requires_grad]
        eng = fused_engine_for(module)
        n_pad = (N + 127) // 128 * 128
        q = x if n_pad == N else torch.cat([x, x.new_zeros(n_pad - N, x.shape[1])], 0)
        _, G = eng.sdf_and_input_grad_rows(q)
        G = G[:N].clone()
        if torch.is_grad_enabled() and deps:
            G = _SecondOrderGuard.apply(G, *deps)
        return None, grad_out * G                   # differentiable w.r.t. grad_out: the jvp trick of deep_sdf/mesh.py:405-422


class _SmallRowsFn(torch.autograd.Function):
    """Eval-mode ``decoder(input[N, L+D])`` of a small network on the fused fp32 small-network kernel (csrc/smallnet.cu): the
    SDF and - when the input requires grad - ``G = dSDF/dinput`` come out of the same launch.  Parameter gradients live on
    ``_ParamBranchFn``; the double-backward trick of ``torch.autograd.functional.jvp`` works as on the other paths."""

    @staticmethod
    def forward(ctx, module, inp):
        from ..small import small_engine_for
        eng = small_engine_for(module)
        x = inp.detach().to(torch.float32).contiguous()
        want_g = inp.requires_grad and torch.is_grad_enabled()
        sdf, G = eng.sdf_and_input_grad(x, with_grad=want_g)
        ctx.module = module
        ctx.have_g = G is not None
        ctx.save_for_backward(x, *([G] if G is not None else []))
        ctx.inp_ref = inp if inp.requires_grad else None
        return sdf.unsqueeze(1)

    @staticmethod
    def backward(ctx, grad_out):
        from ..small import small_engine_for
        module = ctx.module
        if not ctx.needs_input_grad[1]:
            return None, None
        if ctx.have_g:
            x, G = ctx.saved_tensors
        else:
            (x,) = ctx.saved_tensors
            _, G = small_engine_for(module).sdf_and_input_grad(x)
        deps = ([ctx.inp_ref] if ctx.inp_ref is not None else []) + [p for p in module.flat_param_list() if p.requires_grad]
        if torch.is_grad_enabled() and deps:
            G = _SecondOrderGuard.apply(G, *deps)
        return None, grad_out * G


class _ParamBranchFn(torch.autograd.Function):
    """The parameter dependence of the fused eval-mode call as its own graph branch: forward contributes zeros, backward
    yields the parameter gradients from the fp32 generic path (recomputed from the saved input)."""

    @staticmethod
    def forward(ctx, module, x, *params):
        ctx.module = module
        ctx.save_for_backward(x.to(torch.float32).contiguous(), *params)
        return torch.zeros(x.shape[0], 1, device=x.device, dtype=torch.float32)

    @staticmethod
    def backward(ctx, grad_out):
        x, *params = ctx.saved_tensors
        need = ctx.needs_input_grad[2:]
        gp = tuple(None for _ in params)
        if any(need):
            with torch.enable_grad():
                ps = [p.detach().requires_grad_(n) for p, n in zip(params, need)]
                y = _DecoderFn.apply(ctx.module, x, False, 0, *ps)
                want = [p for p in ps if p.requires_grad]
                got = iter(torch.autograd.grad(y, want, grad_out.detach()))
            gp = tuple(next(got) if n else None for n in need)
        return (None, None) + gp


class _DecoderFn(torch.autograd.Function):
    """Generic-path autograd node.  forward/backward are C-ABI calls; for ``create_graph=True`` (the double-backward
    trick behind ``torch.autograd.functional.jvp``, deep_sdf/mesh.py:405-422) backward returns ``grad_out * G`` with the
    per-row input gradient G from the kernel, which is differentiable w.r.t. ``grad_out``."""

    @staticmethod
    def forward(ctx, module: Decoder, inp: torch.Tensor, training: bool, seed: int, *params):
        lib = L.lib()
        desc = module._desc
        x = inp.detach().contiguous().float()
        N = x.shape[0]
        out = torch.empty(N, 1, device=x.device, dtype=torch.float32)
        ctx.module, ctx.training, ctx.seed = module, training, seed
        ctx.n_params = len(params)
        ctx.saved_ws = None
        if N == 0:
            ctx.save_for_backward(x, *params)
            return out
        P = module.c_params([p.detach() for p in params])
        with torch.cuda.device(x.device):
            st = L.stream_ptr(x.device)
            need_grad = any(ctx.needs_input_grad)
            chunk = N
            if lib.dsdf_generic_workspace_bytes(C.byref(desc), N) > _WS_SINGLE_PASS_LIMIT:
                chunk = _CHUNK_ROWS
            ctx.chunk = chunk
            for r0 in range(0, N, chunk):
                n = min(chunk, N - r0)
                nbytes = lib.dsdf_generic_workspace_bytes(C.byref(desc), n)
                ws = torch.empty(nbytes, dtype=torch.uint8, device=x.device)
                L.check(lib.dsdf_generic_forward(C.byref(desc), C.byref(P), x[r0:r0 + n].data_ptr(), n, int(training), seed, r0,
                                                 out[r0:r0 + n].data_ptr(), ws.data_ptr(), nbytes, st))
                if need_grad and chunk == N:
                    ctx.saved_ws = ws
        ctx.save_for_backward(x, *params)
        ctx.inp_ref = inp if inp.requires_grad else None      # only to hang the second-order guard on (backward, create_graph)
        return out

    @staticmethod
    def backward(ctx, grad_out):
        x, *params = ctx.saved_tensors
        module: Decoder = ctx.module
        lib = L.lib()
        desc = module._desc
        N = x.shape[0]
        need_in = ctx.needs_input_grad[1]
        need_p = any(ctx.needs_input_grad[4:])
        none4 = (None, None, None, None)
        if N == 0:
            return none4[:1] + ((torch.zeros_like(x) if need_in else None),) + (None, None) + tuple(
                torch.zeros_like(p) if n else None for p, n in zip(params, ctx.needs_input_grad[4:]))
        double = torch.is_grad_enabled() and grad_out.requires_grad
        P = module.c_params([p.detach() for p in params])
        go = (torch.ones(N, 1, device=x.device) if double else grad_out.detach()).contiguous().float()
        gin = torch.empty_like(x) if (need_in or double) else None
        pgrads = [torch.empty_like(p) for p in params] if (need_p and not double) else None
        if double and need_p:
            # loss.backward(create_graph=True) / a gradient-penalty term: the parameters still get their FIRST-order gradient
            # (one more backward launch with the real cotangent); it is returned as a constant of the higher-order graph
            pgrads2 = _DecoderFn._param_grads_only(ctx, module, x, params, P, grad_out.detach().contiguous().float())
        else:
            pgrads2 = None
        PG = None
        if pgrads is not None:
            PG = L.ParamGrads()
            it = iter(pgrads)
            for layer in range(module.spec.num_linear):
                live = module.layer_tensors(layer)
                for name, t in zip(("v", "g", "b", "ln_w", "ln_b"), live):
                    if t is not None:
                        getattr(PG, name)[layer] = next(it).data_ptr()
        with torch.cuda.device(x.device):
            st = L.stream_ptr(x.device)
            chunk = ctx.chunk
            for ci, r0 in enumerate(range(0, N, chunk)):
                n = min(chunk, N - r0)
                nbytes = lib.dsdf_generic_workspace_bytes(C.byref(desc), n)
                if ctx.saved_ws is not None:
                    ws = ctx.saved_ws
                else:   # recompute this chunk's forward (same Philox key and row ids -> same dropout masks)
                    ws = torch.empty(nbytes, dtype=torch.uint8, device=x.device)
                    scratch = torch.empty(n, 1, device=x.device)
                    L.check(lib.dsdf_generic_forward(C.byref(desc), C.byref(P), x[r0:r0 + n].data_ptr(), n, int(ctx.training),
                                                     ctx.seed, r0, scratch.data_ptr(), ws.data_ptr(), nbytes, st))
                L.check(lib.dsdf_generic_backward(
                    C.byref(desc), C.byref(P), x[r0:r0 + n].data_ptr(), n, int(ctx.training), ctx.seed, r0,
                    go[r0:r0 + n].data_ptr(), gin[r0:r0 + n].data_ptr() if gin is not None else None,
                    C.byref(PG) if PG is not None else None, int(ci > 0), ws.data_ptr(), nbytes, st))
        # create_graph=True: the kernel's input gradient is a constant of the higher-order graph; hang it on a guard node so that
        # a request for ITS derivative (second order in x or in the parameters) raises instead of contributing a silent zero
        deps = [t for t in ([ctx.inp_ref] if ctx.inp_ref is not None else []) + [p for p in params if p.requires_grad]]
        if double:
            # first-order exact; differentiable w.r.t. grad_out (rows are independent scalars)
            G = _SecondOrderGuard.apply(gin, *deps) if deps else gin
            gp = tuple(g if n else None for g, n in zip(pgrads2, ctx.needs_input_grad[4:])) if pgrads2 is not None else tuple(
                None for _ in params)
            return (None, grad_out * G, None, None) + gp
        grads_p = tuple(g if n else None for g, n in zip(pgrads, ctx.needs_input_grad[4:])) if pgrads is not None else tuple(
            None for _ in params)
        if need_in and torch.is_grad_enabled() and deps:
            gin = _SecondOrderGuard.apply(gin, *deps)
        return (None, gin if need_in else None, None, None) + grads_p

    @staticmethod
    def _param_grads_only(ctx, module, x, params, P, go):
        """Parameter gradients for cotangent ``go`` (no input gradient): the backward launches of the normal path."""
        lib = L.lib()
        desc = module._desc
        N = x.shape[0]
        pgrads = [torch.empty_like(p) for p in params]
        PG = L.ParamGrads()
        it = iter(pgrads)
        for layer in range(module.spec.num_linear):
            for name, t in zip(("v", "g", "b", "ln_w", "ln_b"), module.layer_tensors(layer)):
                if t is not None:
                    getattr(PG, name)[layer] = next(it).data_ptr()
        with torch.cuda.device(x.device):
            st = L.stream_ptr(x.device)
            chunk = ctx.chunk
            for ci, r0 in enumerate(range(0, N, chunk)):
                n = min(chunk, N - r0)
                nbytes = lib.dsdf_generic_workspace_bytes(C.byref(desc), n)
                if ctx.saved_ws is not None:
                    ws = ctx.saved_ws
                else:
                    ws = torch.empty(nbytes, dtype=torch.uint8, device=x.device)
                    scratch = torch.empty(n, 1, device=x.device)
                    L.check(lib.dsdf_generic_forward(C.byref(desc), C.byref(P), x[r0:r0 + n].data_ptr(), n, int(ctx.training),
                                                     ctx.seed, r0, scratch.data_ptr(), ws.data_ptr(), nbytes, st))
                L.check(lib.dsdf_generic_backward(C.byref(desc), C.byref(P), x[r0:r0 + n].data_ptr(), n, int(ctx.training), ctx.seed,
                                                  r0, go[r0:r0 + n].data_ptr(), None, C.byref(PG), int(ci > 0), ws.data_ptr(), nbytes, st))
        return pgrads
