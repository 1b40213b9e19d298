"""ctypes binding of libpb_b200.so (the C ABI declared in include/pb_b200.h).

The library is the only compute path: if it cannot be loaded, or the device is not sm_100, every op
raises -- there is no eager/CPU fallback.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libpb_b200.so")

F32, BF16 = 0, 1
ACT_NONE, ACT_GELU, ACT_SIGMOID = 0, 1, 2

_p, _i, _ll, _sz, _f = C.c_void_p, C.c_int, C.c_longlong, C.c_size_t, C.c_float

# name -> argtypes (restype is int unless listed in _RESTYPES)
PROTOTYPES = {
    "pb_version": [],
    "pb_last_error": [],
    "pb_device_ok": [_i],
    "pb_launch_count": [],
    "pb_set_tensor_cores": [_i],
    "pb_cast_f32_to_bf16": [_p, _p, _ll, _p],
    "pb_stage_weights": [_p, _i, _i, _p],
    "pb_flat_copy": [_p, _i, _p, _i, _f, _p],
    "pb_multimem_allreduce": [_p, _ll, _ll, _i, _i, _i, _p],
    "pb_adam_step": [_p, _i, _f, C.c_double, C.c_double, _f, _f, _ll, _p],
    "pb_conv_weight_pack": [_p, _p, _p, _i, _i, _i, _p],
    "pb_embedding_fwd": [_p, _p, _p, _ll, _i, _i, _i, _p],
    "pb_embedding_bwd": [_p, _p, _p, _ll, _i, _i, _i, _p],
    "pb_linear_fwd": [_p, _p, _p, _p, _p, _i, _i, _i, _i, _p, _sz, _p],
    "pb_linear_fwd_noalloc": [_p, _p, _p, _p, _p, _i, _i, _i, _i, _p],
    "pb_linear_bwd": [_p, _p, _p, _p, _p, _p, _p, _p, _i, _i, _i, _i, _i, _p, _sz, _p],
    "pb_linear_tc_ok": [_i, _i, _i],
    "pb_linear_tc_workspace": [_i, _i, _i],
    "pb_linear_fwd_tc": [_p, _p, _p, _p, _p, _i, _i, _i, _i, _p, _sz, _p],
    "pb_linear_bwd_tc": [_p, _p, _p, _p, _p, _p, _p, _p, _i, _i, _i, _i, _i, _p, _sz, _p],
    "pb_res_ln_rows_fwd": [_p, _p, _p, _p, _i, _p, _p, _p, _p, _p, _i, _i, _p],
    "pb_res_ln_rows_bwd": [_p, _p, _p, _p, _p, _p, _i, _p, _p, _p, _p, _p, _i, _i, _p],
    "pb_global_fused_ok": [_i, _i, _i],
    "pb_global_pre_fwd": [_p] * 9 + [_i, _i, _i, _p],
    "pb_global_pre_bwd": [_p] * 12 + [_i, _i, _i, _i, _p, _p],
    "pb_matmul_tn": [_p, _p, _p, _i, _i, _i, _p],
    "pb_global_post_fwd": [_p, _p, _p, _p, _i] + [_p] * 13 + [_i, _i, _p],
    "pb_global_post_bwd": [_p] * 12 + [_i] + [_p] * 12 + [_i, _i, _p],
    "pb_gtrack_ok": [_i, _i, _i],
    "pb_matmul_tn3": [_p, _p, _p, _i, _i, _p, _p, _p, _i, _i, _p, _p, _p, _i, _i, _i, _p],
    "pb_gtrack_fwd": [_p, _p, _p, _p, _i] + [_p] * 22 + [_i, _i, _i, _p],
    "pb_gtrack_bwd": [_p] * 24 + [_i] + [_p] * 13 + [_i, _i, _i, _p],
    "pb_conv_pair_fwd": [_p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _i, _i, _i, _i, _i, _i, _p],
    "pb_conv_pair_bwd": [_p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _p, _i, _i, _i, _i, _i, _i, _p, _sz, _p],
    "pb_ln_joint_stats": [_p, _p, _i, _i, _i, _i, _p],
    "pb_ln_joint_apply": [_p, _p, _p, _p, _p, _i, _i, _i, _f, _i, _p],
    "pb_ln_joint_bwd": [_p, _p, _p, _p, _p, _p, _p, _p, _i, _i, _i, _f, _i, _p],
    "pb_ln_joint_bwd_gelu_ok": [_i, _i, _i],
    "pb_ln_joint_bwd_gelu": [_p] * 15 + [_i, _i, _i, _f, _i, _p],
    "pb_local_linear_fwd": [_p, _p, _p, _p, _p, _p, _i, _i, _i, _i, _p],
    "pb_local_linear_bwd": [_p, _p, _p, _p, _p, _p, _p, _p, _p, _i, _i, _i, _i, _p, _sz, _p],
    "pb_attn_pool_fwd": [_p, _p, _p, _i, _i, _i, _i, _i, _p],
    "pb_attn_pool_bwd": [_p, _p, _p, _p, _p, _p, _p, _p, _i, _i, _i, _i, _i, _i, _p, _sz, _p],
    "pb_attention_smem": [_i, _i, _i, _i, _i],
    "pb_attention_fwd": [_p] * 10 + [_i] * 8 + [_p],
    "pb_attention_bwd": [_p] * 17 + [_i] * 8 + [_p],
    "pb_token_head_fwd": [_p, _p, _p, _p, _i, _i, _i, _i, _i, _p],
    "pb_token_head_bwd": [_p, _p, _p, _p, _p, _p, _p, _p, _i, _i, _i, _i, _i, _p],
    "pb_token_head_tc_ok": [_i, _i, _i, _i],
    "pb_token_head_fwd_tc": [_p, _p, _p, _p, _p, _i, _i, _i, _i, _p],
    "pb_token_head_bwd_tc": [_p, _p, _p, _p, _p, _p, _p, _p, _p, _i, _i, _i, _i, _p, _sz, _p],
    "pb_pretrain_loss": [_p, _p, _p, _p, _p, _p, _p, _p, _p, _i, _i, _i, _i, _p],
    "pb_sigmoid_bwd": [_p, _p, _p, _p, _ll, _p],
    "pb_corrupt_pad": [_p] * 16 + [_i, _i, _i, _p],
    "pb_draw_batch": [C.c_ulonglong, C.c_ulonglong] + [_p] * 8 + [_i, _i, _i, _f, _f, _f, _p],
    "pb_debug_buffer": [_p],
    "pb_workspace_bytes": [C.c_char_p, _i, _i, _i],
    "pb_tc_path": [C.c_char_p, _i, _i, _i, _i, _i, _i],
}
_RESTYPES = {"pb_last_error": C.c_char_p, "pb_launch_count": _ll, "pb_workspace_bytes": _sz, "pb_linear_tc_workspace": _sz, "pb_attention_smem": _sz,
             "pb_debug_buffer": None}

_lib = None


class PBError(RuntimeError):
    pass


def lib():
    """Load (once) and return the ctypes handle.  Raises if the shared library is missing."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise PBError(
                f"{LIB_PATH} is missing: build it with `python __graft_entry__.py build` "
                "(nvcc, sm_100a). There is no fallback path.")
        h = C.CDLL(LIB_PATH)
        for name, args in PROTOTYPES.items():
            fn = getattr(h, name)
            fn.argtypes = args
            fn.restype = _RESTYPES.get(name, _i)
        _lib = h
    return _lib


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = lib().pb_last_error()
        raise PBError(f"{what} failed (code {rc}): {msg.decode() if msg else ''}")


def launch_count() -> int:
    return int(lib().pb_launch_count())
