"""ctypes binding of libm2f_b200.so (the C ABI declared in include/m2f_b200.h).

The library is the only compute path: if it cannot be loaded, or a call fails, a RuntimeError is
raised -- there is no eager / CPU fallback anywhere in this package.
"""
from __future__ import annotations

import ctypes as C
import os
import threading

HERE = os.path.dirname(os.path.abspath(__file__))
# M2F_LIB_PATH: load another build of the same library (A/B of build-time variants, tools/gpu_r2_ab.sh); default: the in-tree .so
LIB_PATH = os.environ.get("M2F_LIB_PATH") or os.path.join(HERE, "libm2f_b200.so")

M2F_OK = 0
M2F_ERR_ARG, M2F_ERR_UNSUPPORTED, M2F_ERR_CUDA, M2F_ERR_ALIGN = -1, -2, -3, -4


class Planes(C.Structure):
    """m2f_planes_t"""
    _fields_ = [("ptr", C.c_void_p), ("rows", C.c_int64), ("cols", C.c_int64), ("ld", C.c_int64),
                ("plane_stride", C.c_int64), ("planes", C.c_int32), ("_pad", C.c_int32)]


class KSeg(C.Structure):
    """m2f_kseg_t"""
    _fields_ = [("a_k0", C.c_int32), ("b_k0", C.c_int32), ("klen", C.c_int32)]


class Epilogue(C.Structure):
    """m2f_epilogue_t"""
    _fields_ = [("bias", C.c_void_p), ("relu", C.c_int32), ("accumulate", C.c_int32),
                ("gate", C.c_void_p), ("gate_ld", C.c_int64), ("gate_scale", C.c_float), ("drop_p", C.c_float),
                ("seed", C.c_uint64), ("step", C.c_uint64), ("site", C.c_uint32), ("_pad", C.c_int32),
                ("drop_ld", C.c_int64), ("resid", C.c_void_p), ("resid_ld", C.c_int64),
                ("out_f32", C.c_void_p), ("out_f32_ld", C.c_int64), ("out_planes", Planes), ("colsum", C.c_void_p)]


class ModalityCfg(C.Structure):
    """m2f_modality_cfg_t"""
    _fields_ = [("enabled", C.c_int32), ("dim", C.c_int32), ("n_head", C.c_int32), ("n_transformers", C.c_int32),
                ("n_layers", C.c_int32)]


class ModelCfg(C.Structure):
    """m2f_model_cfg_t"""
    _fields_ = [("audio", ModalityCfg), ("text", ModalityCfg), ("fam_enabled", C.c_int32), ("fam_dim", C.c_int32),
                ("fam_heads", C.c_int32), ("fam_layers", C.c_int32), ("cls_hidden", C.c_int32), ("cls_out", C.c_int32),
                ("cls_layers", C.c_int32), ("dropout", C.c_float), ("ffn_dim", C.c_int32), ("planes", C.c_int32),
                ("planes_bwd", C.c_int32)]


XCHG_MAX_RANKS, XCHG_MAX_CTAS = 16, 64
XCHG_SIGNAL_BYTES = XCHG_MAX_CTAS * XCHG_MAX_RANKS * 4
XCHG_NO_EXIT_BARRIER = 1
XCHG_WIRE_BF16 = 2


class PeerGroup(C.Structure):
    """m2f_peer_group_t"""
    _fields_ = [("mc_base", C.c_void_p), ("peer_base", C.c_void_p * XCHG_MAX_RANKS), ("signal_off_bytes", C.c_int64),
                ("rank", C.c_int32), ("world", C.c_int32), ("wire_off_bytes", C.c_int64), ("wire_bf16", C.c_int32), ("_pad", C.c_int32)]


BUCKET_CB = C.CFUNCTYPE(None, C.c_void_p, C.c_int64, C.c_int64, C.c_void_p)  # m2f_bucket_cb

PP = C.POINTER(Planes)
_i64, _i32, _u64, _u32, _f32, _vp = C.c_int64, C.c_int32, C.c_uint64, C.c_uint32, C.c_float, C.c_void_p

# name -> argtypes; must list every symbol include/m2f_b200.h declares (tests/test_abi_symbols.py checks)
PROTOTYPES = {
    "m2f_abi_version": [],
    "m2f_last_error": [],
    "m2f_check_device": [],
    "m2f_launch_count": [],
    "m2f_set_rng_source": [C.c_void_p],
    "m2f_gemm": [PP, C.c_int, PP, C.c_int, C.POINTER(KSeg), C.c_int, _i64, _i64, C.POINTER(Epilogue), _vp],
    "m2f_split_planes": [_vp, _i64, PP, _vp],
    "m2f_relu_planes": [PP, PP, _vp],
    "m2f_colsum_planes": [PP, _vp, _vp],
    "m2f_ln_fwd": [_vp, _i64, _vp, _vp, _vp, _i64, _i64, _i64, _vp, _i64, PP, _f32, _u64, _u64, _u32, _vp, _vp],
    "m2f_ln_bwd": [_vp, _i64, _vp, _i64, _vp, _vp, _i64, _i64, _vp, _i64, PP, _f32, _u64, _u64, _u32, _vp, _vp, _vp, _vp],
    "m2f_attn_fwd": [PP, PP, PP, _vp, _i64, _i64, _i32, _f32, _u64, _u64, _u32, PP, _vp],
    "m2f_attn_bwd": [PP, PP, PP, PP, _vp, _i64, _i64, _i32, _f32, _u64, _u64, _u32, PP, PP, PP, _vp, _vp, _vp, _vp],
    "m2f_ce_fwd_bwd": [_vp, _i64, _vp, _vp, _i64, _i32, _f32, _f32, _vp, _vp, _i64, PP, _vp],
    "m2f_ce_fwd_bwd_static": [_vp, _i64, _vp, _vp, _i64, _i32, _f32, _f32, _vp, _vp, _i64, PP, _vp],
    "m2f_dropout_mask": [_f32, _u64, _u64, _u32, _i64, _vp, _vp],
    "m2f_model_create": [C.POINTER(ModelCfg), C.POINTER(_vp)],
    "m2f_model_destroy": [_vp],
    "m2f_model_num_params": [_vp],
    "m2f_model_param_info": [_vp, _i64, C.c_char_p, _i32, C.POINTER(_i64), C.POINTER(_i64), C.POINTER(_i64)],
    "m2f_model_arena_floats": [_vp],
    "m2f_model_arena_1d_offset": [_vp],
    "m2f_model_shadow_bytes": [_vp],
    "m2f_model_refresh_shadow": [_vp, _vp, _vp, _vp],
    "m2f_plan_create": [_vp, _i64, _i64, _i32, C.POINTER(_vp)],
    "m2f_plan_destroy": [_vp],
    "m2f_plan_workspace_bytes": [_vp],
    "m2f_plan_forward": [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _i32, _u64, _u64, _vp, _vp],
    "m2f_plan_loss": [_vp, _vp, _vp, _vp, _vp, _f32, _f32, _vp, _vp],
    "m2f_plan_backward": [_vp, _vp, _vp, _vp, _vp, _vp, _vp],
    "m2f_plan_backward_bucketed": [_vp, _vp, _vp, _vp, _vp, _vp, _i64, _vp, _vp, _vp],
    "m2f_grad_allreduce_peer": [C.POINTER(PeerGroup), _i64, _i64, _f32, _i32, _i32, _vp],
    "m2f_exchange_trace": [_vp],
    "m2f_exchange_tune": [_i32, _i32],
    "m2f_plan_set_exchange": [_vp, C.POINTER(PeerGroup), _f32, _i32],
    "m2f_plan_exchange_ranges": [_vp],
    "m2f_plan_set_optimizer": [_vp, _vp, _vp, _vp],
    "m2f_adam_hyper": [_vp, _f32, _f32, _f32, _f32, _f32, _i64, _f32, _vp],
    "m2f_adam_step": [_vp, _vp, _vp, _vp, _i64, _f32, _f32, _f32, _f32, _f32, _i64, _f32, _vp],
    "m2f_model_adam_step": [_vp, _vp, _vp, _vp, _vp, _vp, _f32, _f32, _f32, _f32, _f32, _i64, _f32, _vp],
    "m2f_collate_dialogues": [_vp, _i64, _i32, _vp, _i64, _i32, _vp, _vp, _vp, _vp, _i64, _i64, _vp, _vp, _vp, _vp, _vp],
    "m2f_batch_metrics": [_vp, _i64, _vp, _i64, _i32, _i64, _vp, _vp, _vp, _vp],
}

# functions that return a 64-bit count instead of an error code
_RET_I64 = {"m2f_launch_count", "m2f_exchange_trace", "m2f_plan_exchange_ranges", "m2f_model_num_params", "m2f_model_arena_floats", "m2f_model_arena_1d_offset", "m2f_model_shadow_bytes",
            "m2f_plan_workspace_bytes"}

_lock = threading.Lock()
_lib = None


def lib() -> C.CDLL:
    """Load (once) and return the shared library; raises if it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is None:
            if not os.path.exists(LIB_PATH):
                raise RuntimeError(
                    f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                    "(nvcc, sm_100a).  This package has no CPU or eager fallback.")
            L = C.CDLL(LIB_PATH)
            for name, args in PROTOTYPES.items():
                fn = getattr(L, name)  # AttributeError if the symbol is missing
                fn.argtypes = args
                fn.restype = C.c_char_p if name == "m2f_last_error" else (C.c_int64 if name in _RET_I64 else C.c_int)
            _lib = L
    return _lib


def check(rc: int, what: str = "") -> None:
    if rc != M2F_OK:
        msg = lib().m2f_last_error()
        raise RuntimeError(f"libm2f_b200 {what} failed (code {rc}): {msg.decode() if msg else '?'}")
