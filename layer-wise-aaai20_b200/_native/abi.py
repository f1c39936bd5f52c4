"""ctypes signatures mirroring ``include/lwc_b200.h``."""
from __future__ import annotations

import ctypes as C

c_void_pp = C.POINTER(C.c_void_p)


class LwcPlanDesc(C.Structure):
    _fields_ = [
        ("num_tensors", C.c_int32), ("num_groups", C.c_int32), ("method", C.c_int32),
        ("grad_dtype", C.c_int32), ("out_dtype", C.c_int32), ("world", C.c_int32),
        ("rank", C.c_int32), ("qstates", C.c_int32), ("threshold", C.c_float),
        ("seed", C.c_uint64),
        ("grad_dev", c_void_pp), ("out_dev", c_void_pp), ("resid_dev", c_void_pp),
        ("numel", C.POINTER(C.c_int64)), ("group_of", C.POINTER(C.c_int32)),
        ("uid", C.POINTER(C.c_int32)),
        ("keep", C.POINTER(C.c_int64)), ("cap", C.POINTER(C.c_int64)),
    ]


class LwcSgdDesc(C.Structure):
    _fields_ = [
        ("num_tensors", C.c_int32), ("grad_dtype", C.c_int32), ("model_dtype", C.c_int32),
        ("lr", C.c_float), ("momentum", C.c_float), ("dampening", C.c_float),
        ("weight_decay", C.c_float), ("grad_scale", C.c_float),
        ("nesterov", C.c_int32), ("first_step", C.c_int32), ("lr_dev", C.c_void_p), ("param_dtype", C.c_int32),
        ("skip_flag", C.c_void_p),
    ]


class LwcSgdTensor(C.Structure):
    _fields_ = [("param", C.c_void_p), ("grad", C.c_void_p), ("mom", C.c_void_p),
                ("model", C.c_void_p), ("n", C.c_int64)]


LWC_F32, LWC_BF16, LWC_F16 = 0, 1, 2
METHOD_CODES = {"none": 0, "Topk": 1, "Randomk": 2, "Thresholdv": 3, "AdaptiveThreshold": 4,
                "TernGrad": 5, "RandomDithering": 6}
LWC_CHUNK = 8192
LWC_MAX_WORLD = 8


def dtype_code(dtype) -> int:
    import torch
    return {torch.float32: LWC_F32, torch.bfloat16: LWC_BF16, torch.float16: LWC_F16}[dtype]


def declare(lib: C.CDLL):
    def sig(name, restype, *argtypes):
        fn = getattr(lib, name, None)
        if fn is None:
            return
        fn.restype = restype
        fn.argtypes = list(argtypes)

    vp, i32, i64, u64, sz = C.c_void_p, C.c_int32, C.c_int64, C.c_uint64, C.c_size_t
    sig("lwc_malloc", C.c_int, c_void_pp, sz)
    sig("lwc_free", C.c_int, vp)
    sig("lwc_memset_async", C.c_int, vp, C.c_int, sz, vp)
    sig("lwc_ipc_get_handle", C.c_int, vp, vp)
    sig("lwc_ipc_base_offset", C.c_int, vp, C.POINTER(C.c_uint64))
    sig("lwc_ipc_open_handle", C.c_int, vp, c_void_pp)
    sig("lwc_ipc_close_handle", C.c_int, vp)
    sig("lwc_enable_peer_access", C.c_int, C.c_int)
    sig("lwc_error_string", C.c_char_p, C.c_int)
    sig("lwc_set_flag_timeout_ms", C.c_int, C.c_uint)
    sig("lwc_plan_create", C.c_int, c_void_pp, C.POINTER(LwcPlanDesc))
    sig("lwc_plan_destroy", C.c_int, vp)
    sig("lwc_plan_slot_bytes", sz, vp)
    sig("lwc_plan_region_bytes", sz, vp)
    sig("lwc_plan_bind", C.c_int, vp, c_void_pp, c_void_pp)
    sig("lwc_plan_use_device_step", C.c_int, vp, C.c_int, u64)
    sig("lwc_plan_bind_multicast", C.c_int, vp, vp)
    sig("lwc_plan_set_random", C.c_int, vp, c_void_pp)
    sig("lwc_plan_set_values_only", C.c_int, vp, C.c_int)
    sig("lwc_plan_update_grads", C.c_int, vp, c_void_pp, vp)
    sig("lwc_plan_send", C.c_int, vp, u64, vp)
    sig("lwc_plan_recv", C.c_int, vp, u64, vp)
    sig("lwc_plan_stats", C.c_int, vp, C.POINTER(u64), vp)
    sig("lwc_dense_allreduce", C.c_int, c_void_pp, c_void_pp, vp, i64, i64, C.c_int, C.c_int, C.c_int, u64, vp)
    sig("lwc_topk_workspace_bytes", sz, i64)
    sig("lwc_topk_select_f32", C.c_int, vp, i64, i64, i64, vp, vp, vp, vp, vp)
    sig("lwc_topk_select_f32_noalloc", C.c_int, vp, i64, i64, i64, vp, vp, vp, vp)
    sig("lwc_gemm_workspace_bytes", sz, C.c_int, C.c_int, C.c_int)
    sig("lwc_gemm_pick_splits", C.c_int, C.c_int, C.c_int, C.c_int)
    sig("lwc_gemm_bf16", C.c_int, vp, vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
        C.c_int, C.c_int, vp, C.c_int, C.c_int, C.c_int, vp, vp)
    sig("lwc_bn_workspace_bytes", sz, C.c_longlong, C.c_int)
    sig("lwc_bn_forward", C.c_int, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, C.c_longlong, C.c_int, C.c_int, C.c_int,
        C.c_int, C.c_float, C.c_float, vp)
    sig("lwc_bn_forward_ex", C.c_int, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, C.c_longlong, C.c_int, C.c_int, C.c_int,
        C.c_int, C.c_float, C.c_float, vp, vp, vp)
    sig("lwc_bn_backward_ex", C.c_int, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, C.c_longlong, C.c_int, C.c_int,
        C.c_int, vp, vp)
    sig("lwc_bn_backward", C.c_int, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, C.c_longlong, C.c_int, C.c_int,
        C.c_int, vp)
    sig("lwc_maxpool_forward", C.c_int, vp, vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
        C.c_int, vp)
    sig("lwc_maxpool_backward", C.c_int, vp, vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
        C.c_int, vp)
    sig("lwc_normalize_u8", C.c_int, vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, vp, vp, C.c_int, vp)
    sig("lwc_cross_entropy_forward", C.c_int, vp, vp, vp, vp, C.c_int, C.c_int, C.c_int, vp)
    sig("lwc_cross_entropy_backward", C.c_int, vp, vp, vp, vp, vp, C.c_int, C.c_int, C.c_int, C.c_float, vp)
    sig("lwc_conv_tiles_m", C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int))
    sig("lwc_conv_forward", C.c_int, vp, vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, vp,
        C.c_int, vp, vp)
    sig("lwc_conv1x1_add", C.c_int, vp, vp, vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, vp)
    sig("lwc_conv1x1_add_mask", C.c_int, vp, vp, vp, vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, vp)
    ci, ll, cf = C.c_int, C.c_longlong, C.c_float
    sig("lwc_convx_max_partials", ci)
    sig("lwc_convx_forward", ci, vp, vp, vp, ci, ci, ci, ci, ci, ci, ci, ci, vp, ci, vp, C.POINTER(ci), ci, vp)
    sig("lwc_convx_dgrad1x1", ci, vp, vp, vp, ll, ci, ci, ci, vp)
    sig("lwc_convx_dgrad", ci, vp, vp, vp, ci, ci, ci, ci, ci, ci, ci, ci, ci, vp)
    sig("lwc_convx_stem_forward", ci, vp, vp, vp, ci, ci, ci, ci, vp, C.POINTER(ci), ci, vp)
    sig("lwc_convx_wgrad_workspace_bytes", sz, ci, ci, ci, ci, ci, ci, ci, ci, ci)
    sig("lwc_convx_wgrad", ci, vp, vp, vp, vp, ci, ci, ci, ci, ci, ci, ci, ci, ci, ci, cf, ci, vp)
    sig("lwc_convx_dgrad_bnstats", ci, vp, vp, vp, ci, ci, ci, ci, ci, ci, ci, vp, vp, vp, vp, vp, ci, vp, C.POINTER(ci), ci, vp)
    sig("lwc_convx_dgrad1x1_fork", ci, vp, vp, vp, vp, vp, ll, ci, ci, vp, vp, vp, vp, C.POINTER(ci), ci, vp)
    sig("lwc_bn_backward_from_partials", ci, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, ci, ll, ci, ci, ci, vp)
    sig("lwc_convx_stats_finalize", ci, vp, ci, ci, ll, cf, cf, vp, vp, vp, vp, vp, vp)
    sig("lwc_bn_apply_stats", ci, vp, vp, vp, vp, vp, vp, vp, ll, ci, ci, ci, vp, vp)
    sig("lwc_eltwise", ci, vp, vp, vp, ll, ci, cf, ci, vp)
    sig("lwc_dropout_forward", ci, vp, vp, vp, ll, cf, u64, u64, ci, vp)
    sig("lwc_dropout_backward", ci, vp, vp, vp, ll, cf, ci, vp)
    sig("lwc_adaptive_avgpool", ci, vp, vp, ci, ci, ci, ci, ci, ci, ci, ci, vp)
    sig("lwc_accuracy", ci, vp, vp, vp, vp, ci, ci, ci, vp)
    sig("lwc_stem_wgrad_workspace_bytes", C.c_size_t, C.c_int, C.c_int)
    sig("lwc_stem_wgrad", C.c_int, vp, vp, vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, vp)
    sig("lwc_sgd_plan_create", C.c_int, c_void_pp, C.POINTER(LwcSgdTensor), i32)
    sig("lwc_sgd_plan_destroy", C.c_int, vp)
    sig("lwc_sgd_step", C.c_int, vp, C.POINTER(LwcSgdDesc), vp)
    sig("lwc_peer_broadcast", C.c_int, c_void_pp, c_void_pp, vp, C.c_longlong, C.c_longlong, C.c_int, C.c_int, C.c_int, u64, vp)
    sig("lwc_poison_ptr", C.c_int, c_void_pp)
    sig("lwc_poison_read", C.c_int, C.POINTER(C.c_uint))
    sig("lwc_poison_clear", C.c_int)


def check(rc: int, what: str = ""):
    if rc != 0:
        from .lib import lib
        msg = lib().lwc_error_string(rc)
        raise RuntimeError(f"{what or 'native call'} failed ({rc}): {msg.decode() if msg else '?'}")
