"""ctypes binding of the C-ABI library ``libdsdf_b200.so`` (see ``include/dsdf.h``).

The library is the product: there is no Python/PyTorch fallback.  If the shared object is
missing, or the current device is not sm_100, every op raises.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional, Sequence

import torch

MAX_LAYERS = 16
ABI_VERSION = 3          # include/dsdf.h DSDF_ABI_VERSION
_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("DSDF_B200_LIB") or os.path.join(_HERE, "libdsdf_b200.so")   # the override is for A/B timing of two builds


class NetDesc(C.Structure):
    _fields_ = [
        ("latent_size", C.c_int32), ("geom_dim", C.c_int32), ("num_linear", C.c_int32),
        ("in_dim", C.c_int32 * MAX_LAYERS), ("out_dim", C.c_int32 * MAX_LAYERS),
        ("weight_norm_mask", C.c_uint32), ("layernorm_mask", C.c_uint32), ("dropout_mask", C.c_uint32),
        ("latent_in_mask", C.c_uint32), ("xyz_in_all", C.c_int32), ("use_tanh", C.c_int32),
        ("latent_dropout", C.c_int32), ("dropout_prob", C.c_float),
    ]


class Params(C.Structure):
    _fields_ = [(n, C.c_void_p * MAX_LAYERS) for n in ("v", "g", "b", "ln_w", "ln_b")]


class ParamGrads(C.Structure):
    _fields_ = [(n, C.c_void_p * MAX_LAYERS) for n in ("v", "g", "b", "ln_w", "ln_b")]


class TrainIO(C.Structure):
    _fields_ = [
        ("b_last", C.c_void_p), ("training", C.c_int32), ("seed", C.c_uint64), ("row0", C.c_uint64),
        ("gt", C.c_void_p), ("gt_stride", C.c_int32), ("clamp", C.c_float), ("inv_n", C.c_float), ("dy_out", C.c_void_p),
        ("tile_loss", C.c_void_p), ("bitmask", C.c_void_p), ("acts", C.c_void_p), ("n_rows_pad", C.c_int64),
        ("dropout_off", C.c_int32), ("dy_unit", C.c_int32), ("tile_rows", C.c_void_p), ("step_state", C.c_void_p),
    ]


class DsdfError(RuntimeError):
    pass


_lib: Optional[C.CDLL] = None


def _declare(lib: C.CDLL) -> None:
    vp, i32, i64, u32, u64, f32, sz = C.c_void_p, C.c_int32, C.c_int64, C.c_uint32, C.c_uint64, C.c_float, C.c_size_t
    P = C.POINTER
    sigs = {
        "dsdf_abi_version": (C.c_int, []),
        "dsdf_last_error": (C.c_char_p, []),
        "dsdf_check_device": (C.c_int, []),
        "dsdf_launch_count": (i64, []),
        "dsdf_dropout_effective_p": (f32, [f32]),
        "dsdf_dropout_scale": (f32, [f32]),
        "dsdf_dropout_mask": (C.c_int, [u64, u32, u64, i64, i32, f32, vp, vp]),
        "dsdf_generic_workspace_bytes": (sz, [P(NetDesc), i64]),
        "dsdf_generic_forward": (C.c_int, [P(NetDesc), P(Params), vp, i64, i32, u64, u64, vp, vp, sz, vp]),
        "dsdf_generic_backward": (C.c_int, [P(NetDesc), P(Params), vp, i64, i32, u64, u64, vp, vp, P(ParamGrads), i32,
                                            vp, sz, vp]),
        "dsdf_small_supported": (C.c_int, [P(NetDesc)]),
        "dsdf_small_pack_bytes": (sz, [P(NetDesc)]),
        "dsdf_small_pack_weights": (C.c_int, [P(NetDesc), P(Params), vp, vp]),
        "dsdf_small_forward": (C.c_int, [P(NetDesc), vp, vp, i64, vp, i64, vp, vp, i64, vp]),
        "dsdf_small_latent_step": (C.c_int, [P(NetDesc), vp, vp, i64, vp, i32, i64, f32, f32, vp, vp, vp, vp, vp, vp]),
        "dsdf_fused_supported": (C.c_int, [P(NetDesc)]),
        "dsdf_fused_rows_supported": (C.c_int, [P(NetDesc)]),
        "dsdf_pack_bytes": (sz, [P(NetDesc)]),
        "dsdf_pack_weights": (C.c_int, [P(NetDesc), P(Params), vp, vp]),
        "dsdf_pack_weights_ex": (C.c_int, [P(NetDesc), P(Params), vp, u32, vp]),
        "dsdf_scene_aux_bytes": (sz, [i32]),
        "dsdf_scene_prepare": (C.c_int, [P(NetDesc), vp, vp, vp, i32, f32, vp, vp]),
        "dsdf_fused_forward": (C.c_int, [P(NetDesc), vp, vp, vp, vp, i32, i64, vp, P(TrainIO), vp]),
        "dsdf_fused_forward_rows": (C.c_int, [P(NetDesc), vp, vp, i32, i64, vp, P(TrainIO), vp]),
        "dsdf_grid_points": (C.c_int, [i32, i64, i64, vp, vp]),
        "dsdf_sample_batch": (C.c_int, [vp, vp, vp, i32, i32, u64, vp, vp]),
        "dsdf_sample_perm": (u32, [u32, u32, u64, i64, i32]),
        "dsdf_tile_queries": (C.c_int, [vp, i32, vp, i32, i32, i64, i64, P(f32), vp, vp]),
        "dsdf_cap_sdf": (C.c_int, [vp, vp, i64, i32, vp, i32, i64, i32, P(i32), P(C.c_double), P(C.c_double), P(i32), vp]),
        "dsdf_iso_num_blocks": (i64, [i32, i32, i32]),
        "dsdf_iso_count": (C.c_int, [vp, i32, i32, i32, f32, vp, vp]),
        "dsdf_iso_emit": (C.c_int, [vp, i32, i32, i32, f32, P(f32), P(f32), vp, vp, vp]),
        "dsdf_exact_pack_bytes": (C.c_size_t, [vp]),
        "dsdf_exact_pack": (C.c_int, [vp, vp, vp, vp]),
        "dsdf_exact_scene_aux": (C.c_int, [vp, vp, vp, vp, i32, vp, vp]),
        "dsdf_fused_forward_exact": (C.c_int, [vp, vp, vp, vp, vp, vp, i32, i64, vp, vp, vp]),
        "dsdf_adam": (C.c_int, [vp, vp, vp, vp, i64, f32, f32, f32, f32, i32, f32, vp, vp]),
        "dsdf_clip_coef": (C.c_int, [vp, i64, f32, vp, vp, vp]),
        "dsdf_nvls_signal_words": (C.c_int32, [i32]),
        "dsdf_nvls_allreduce_f32": (C.c_int, [vp, vp, i32, i32, i64, vp]),
        "dsdf_clip_coef_static": (C.c_int, [vp, i64, f32, vp, vp]),
        "dsdf_adam_state": (C.c_int, [vp, vp, vp, vp, i64, f32, f32, f32, f32, vp, f32, vp, vp]),
        "dsdf_step_advance": (C.c_int, [vp, f32, f32, vp]),
        "dsdf_step_set": (C.c_int, [vp, u32, vp]),
        "dsdf_selftest_mma_rate": (C.c_int, [C.c_int, C.c_int, C.c_int, vp, vp]),
        "dsdf_selftest_umma_pair": (C.c_int, [C.c_int, vp, vp, vp, C.c_int, vp]),
        "dsdf_selftest_umma_aux": (C.c_int, [vp, vp, vp, vp, vp, C.c_int, C.c_int, C.c_uint, C.c_uint, vp]),
        "dsdf_selftest_umma": (C.c_int, [C.c_int, vp, vp, vp, vp, C.c_int, C.c_int, C.c_uint, C.c_uint, C.c_uint,
                                         C.c_uint, vp]),
    }
    for name, (res, args) in sigs.items():
        if hasattr(lib, name):
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args


def lib() -> C.CDLL:
    """Load (once) and return the C-ABI library.  Raises if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise DsdfError(
                f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "or `make -C deepsdf_b200/csrc` (there is no fallback path)")
        _lib = C.CDLL(LIB_PATH)
        _declare(_lib)
        if _lib.dsdf_abi_version() != ABI_VERSION:
            raise DsdfError("libdsdf_b200.so ABI version mismatch")
    return _lib


def check(rc: int) -> None:
    if rc != 0:
        msg = lib().dsdf_last_error()
        raise DsdfError(f"dsdf error {rc}: {msg.decode() if msg else '?'}")


def ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    if t is None:
        return None
    return t.data_ptr()


def stream_ptr(device=None) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def require_cuda(*tensors: torch.Tensor) -> None:
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise DsdfError("deepsdf_b200 ops need CUDA tensors (sm_100a); there is no CPU path")
