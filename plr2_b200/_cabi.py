"""ctypes binding of the C-ABI in include/plr2_b200.h.

There is deliberately NO fallback: if libplr2_b200.so is missing or a call fails, we raise.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libplr2_b200.so")

_vp, _i, _f, _sz = C.c_void_p, C.c_int, C.c_float, C.c_size_t

# name -> (restype, argtypes); must list every symbol include/plr2_b200.h declares
SIGNATURES = {
    "plr2_abi_version": (_i, []),
    "plr2_last_error_string": (C.c_char_p, []),
    "plr2_device_sm_count": (_i, []),
    "plr2_launch_count": (C.c_longlong, []),
    "plr2_profile_begin": (_i, []),
    "plr2_profile_end": (_i, [C.POINTER(C.c_double), C.POINTER(C.c_longlong), C.POINTER(C.c_double)]),
    "plr2_profile_executed_flops": (C.c_double, []),
    "plr2_knn_workspace_bytes": (_sz, [_i] * 5),
    "plr2_knn_f32": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _vp, _sz, _vp]),
    "plr2_knn_plan": (_i, [_i, _i, _i, C.POINTER(_i), C.POINTER(_i)]),
    "plr2_pose_loss_fwd": (_i, [_vp] * 8 + [_i] * 4 + [_f, _i, _i] + [_vp] * 5),
    "plr2_pose_loss_bwd": (_i, [_vp] * 7 + [_i, _i, _f] + [_vp] * 4),
    "plr2_pose_select": (_i, [_vp] * 6 + [_i] * 3 + [_vp] * 6),
    "plr2_pose_compose": (_i, [_vp, _vp, _i, _vp]),
    "plr2_set_backward_stage_event": (_i, [_vp]),
    "plr2_net_grad_regions": (_i, [_i, _i, C.POINTER(_sz), C.POINTER(_sz)]),
    "plr2_pose_metric_workspace_bytes": (_sz, [_i]),
    "plr2_pose_metric": (_i, [_vp, C.c_longlong, _vp, _vp, _i, _i, _vp, _vp, _vp, _sz, _vp]),
    "plr2_mask_bbox": (_i, [_vp, _i, _vp, _vp, _i, _i, _i, _vp, _vp, _vp, _vp]),
    "plr2_choose_workspace_bytes": (_sz, [_i, _i]),
    "plr2_choose_from_mask": (_i, [_vp, _i, _i, _vp, _i, _i, _vp, C.c_longlong, C.c_ulonglong, _vp, _vp, _vp, _sz, _vp]),
    "plr2_backproject": (_i, [_vp, _i, _i, _i, _vp, _vp, _vp, _vp, _i, _i, _vp, _vp]),
    "plr2_transform_target": (_i, [_vp, _vp, _vp, _vp, _i, _i, _vp, _vp]),
    "plr2_pose_errors": (_i, [_vp, _vp, _i, _vp, _vp, _vp]),
    "plr2_auc_workspace_bytes": (_sz, [_i]),
    "plr2_auc_f32": (_i, [_vp, _i, C.c_double, C.c_double, _vp, _vp, _sz, _vp]),
    "plr2_refine_loss_workspace_bytes": (_sz, [_i, _i]),
    "plr2_refine_loss_fwd": (_i, [_vp] * 7 + [_i] * 5 + [_vp] * 6 + [_vp, _sz, _vp]),
    "plr2_kalman_filter_f64": (_i, [_vp] * 6 + [C.c_double, C.c_double, _i, _i] + [_vp] * 4),
    "plr2_refine_loss_static_scratch_bytes": (_sz, []),
    "plr2_refine_loss_fwd_static": (_i, [_vp] * 7 + [_i] * 5 + [_vp] * 6 + [_vp]),
}



class GemmDesc(C.Structure):
    _fields_ = [("A", _vp), ("sam", C.c_longlong), ("sak", C.c_longlong),
                ("B", _vp), ("sbk", C.c_longlong), ("sbn", C.c_longlong),
                ("C", _vp), ("ldc", C.c_longlong),
                ("M", _i), ("N", _i), ("K", _i),
                ("bias", _vp), ("frame_bias", _vp), ("rows_per_frame", _i), ("ldfb", C.c_longlong),
                ("relu_mask", _vp), ("ldmask", C.c_longlong),
                ("relu", _i), ("accumulate", _i), ("split_k", _i)]


POSENET_LAYERS = 18
REFINER_LAYERS = 12


class ParamTable(C.Structure):
    _fields_ = [("w", _vp * POSENET_LAYERS), ("b", _vp * POSENET_LAYERS)]


_pp = C.POINTER(ParamTable)
SIGNATURES.update({
    "plr2_gemm_f32": (_i, [C.POINTER(GemmDesc), _vp]),
    "plr2_split_bf16x2": (_i, [_vp, C.c_longlong, C.c_longlong, _i, _vp, C.c_longlong, C.c_longlong, _vp]),
    "plr2_gemm_bf16x2": (_i, [_vp, C.c_longlong, C.c_longlong, _vp, C.c_longlong, C.c_longlong, _i, _i, _i, _vp, _i, _vp, C.c_longlong, _vp]),
    "plr2_gemm_bf16x2_tn": (_i, [_vp, C.c_longlong, C.c_longlong, _vp, C.c_longlong, C.c_longlong, _i, _i, _i, _vp, C.c_longlong, _vp]),
    "plr2_net_packed_floats": (_sz, [_i, _i]),
    "plr2_net_input_status": (_i, [_vp, C.POINTER(C.c_uint), _vp]),
    "plr2_net_workspace_bytes": (_sz, [_i] * 5),
    "plr2_net_pack": (_i, [_i, _pp, _i, _vp, _i, _vp]),
    "plr2_net_unpack_grads": (_i, [_i, _vp, _pp, _i, _vp]),
    "plr2_net_accumulate_grads": (_i, [_i, _vp, _pp, _i, _vp]),
    "plr2_posenet_forward": (_i, [_vp] * 5 + [_i] * 4 + [_vp, _sz] + [_vp] * 4 + [_i, _i, _vp]),
    "plr2_posenet_backward": (_i, [_vp] * 4 + [_i] * 4 + [_vp, _sz] + [_vp] * 5 + [_i, _vp]),
    "plr2_refiner_forward": (_i, [_vp] * 4 + [_i] * 3 + [_vp, _sz] + [_vp] * 2 + [_i, _i, _vp]),
    "plr2_refiner_backward": (_i, [_vp] * 4 + [_i] * 3 + [_vp, _sz] + [_vp] * 3 + [_i, _vp]),
})

_lib = None


class Plr2Error(RuntimeError):
    pass


def lib():
    """Load (once) and return the ctypes handle. Raises if the CUDA library was not built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise Plr2Error(
                f"{LIB_PATH} not found: build it with `python -m plr2_b200.build` "
                "(there is no CPU / PyTorch fallback for this path)")
        h = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(h, name)  # AttributeError if the symbol is missing
            fn.restype = res
            fn.argtypes = args
        _lib = h
    return _lib


def check(status, what):
    if status != 0:
        msg = lib().plr2_last_error_string().decode(errors="replace")
        raise Plr2Error(f"{what} failed with status {status}: {msg}")


def ptr(t):
    """Device/host pointer of a contiguous torch tensor (None -> NULL)."""
    if t is None:
        return None
    assert t.is_contiguous(), "plr2_b200 C-ABI needs contiguous tensors"
    return t.data_ptr()


def stream_ptr(device=None):
    import torch
    return torch.cuda.current_stream(device).cuda_stream
