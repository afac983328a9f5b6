"""ctypes binding of libwxddp.so (include/wxddp.h).  There is no fallback: if the library is missing or
a call fails, the caller gets an exception."""
from __future__ import annotations

import ctypes as C
import os

from . import _build

WX_F32, WX_F64 = 0, 1
IMPUTE_NONE, IMPUTE_MEAN = 0, 1
DROPOUT_OFF, DROPOUT_MASK, DROPOUT_PHILOX = 0, 1, 2
MAX_FEATURES = 32


class ModelDesc(C.Structure):
    _fields_ = [("n_features", C.c_int32), ("hidden", C.c_int32), ("classes", C.c_int32),
                ("dropout_p", C.c_float), ("lr", C.c_double), ("beta1", C.c_double),
                ("beta2", C.c_double), ("eps", C.c_double), ("optimizer", C.c_int32)]


OPT_ADAM, OPT_SGD = 0, 1

class WideDesc(C.Structure):
    _fields_ = [("n_features", C.c_int32), ("hidden", C.c_int32), ("n_hidden", C.c_int32), ("classes", C.c_int32),
                ("dropout_p", C.c_float), ("lr", C.c_double), ("beta1", C.c_double), ("beta2", C.c_double),
                ("eps", C.c_double)]


class WxError(RuntimeError):
    pass


_vp = C.c_void_p
_i64 = C.c_int64
_u64 = C.c_uint64
_int = C.c_int

# name -> (restype, argtypes); every symbol declared in include/wxddp.h
SIGNATURES = {
    "wx_version": (_int, []),
    "wx_build_arch": (C.c_char_p, []),
    "wx_last_error": (C.c_char_p, []),
    "wx_create": (_int, [_int, C.POINTER(ModelDesc), C.POINTER(_vp)]),
    "wx_destroy": (_int, [_vp]),
    "wx_n_params": (_i64, [_vp]),
    "wx_num_sms": (_int, [_vp]),
    "wx_launch_count": (_i64, [_vp]),
    "wx_peer_export": (_int, [_vp, _vp]),
    "wx_peer_connect": (_int, [_vp, _int, _int, _vp]),
    "wx_label_encode": (_int, [_vp, _vp, _vp, _vp, _i64, _vp, _vp, _vp]),
    "wx_col_stats": (_int, [_vp, C.POINTER(_vp), _int, C.POINTER(_vp), _i64, _vp, _vp]),
    "wx_stats_merge": (_int, [_vp, _vp, _int, _vp, _vp]),
    "wx_normalize_pack": (_int, [_vp, C.POINTER(_vp), _int, C.POINTER(_vp), _i64, _vp, _int, _vp, _vp, _vp]),
    "wx_zscore_f64": (_int, [_vp, C.POINTER(_vp), _int, C.POINTER(_vp), _i64, _vp, C.POINTER(_vp), _vp]),
    "wx_impute_index_list": (_int, [_vp, _vp, _i64, _vp, _i64, _vp, _vp]),
    "wx_mlp_grad": (_int, [_vp, _vp, _vp, _vp, _vp, _i64, _i64, _int, _vp, _u64, _u64, C.c_float, _vp, _vp]),
    "wx_mlp_grad_static": (_int, [_int, _vp, _vp, _vp, _vp, _i64, _i64, _int, _vp, _u64, _u64, C.c_float, _vp, _vp]),
    "wx_tc_mlp_grad": (_int, [_vp, _vp, _vp, _vp, _i64, _i64, _int, _vp, _u64, _u64, C.c_float, _vp, _vp]),
    "wx_tc_train_step": (_int, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _i64, _i64, _int, _vp, _u64, _vp, _vp]),
    "wx_tc_train_step_raw": (_int, [_vp, _vp, _vp, _vp, _vp, C.POINTER(_vp), _vp, _int, _vp, _i64, _i64, _int, _vp, _u64,
                                    _vp, _vp]),
    "wx_tc_last_status": (_int, [_vp, C.POINTER(_int)]),
    "wx_adam_step": (_int, [_vp, _vp, _vp, _vp, _vp, _i64, _i64, _vp]),
    "wx_train_step": (_int, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i64, _i64, _int, _vp, _u64, _vp, _vp]),
    "wx_train_step_raw": (_int, [_vp, _vp, _vp, _vp, _vp, C.POINTER(_vp), _int, _vp, _int, _vp, _i64, _i64, _int,
                                 _vp, _u64, _vp, _vp]),
    "wx_train_steps_small": (_int, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i64, _i64, _i64, _int, _vp, _u64,
                                    _vp, _vp]),
    "wx_mlp_eval": (_int, [_vp, _vp, _vp, _vp, _vp, _i64, _i64, _vp, _vp]),
    "wx_mlp_predict": (_int, [_vp, _vp, _vp, _i64, _vp, _vp]),
    "wx_mlp_logits": (_int, [_vp, _vp, _vp, _i64, _vp, _vp]),
    "wx_wide_n_params": (_i64, [C.POINTER(WideDesc)]),
    "wx_wide_workspace_bytes": (_i64, [C.POINTER(WideDesc), _i64]),
    "wx_wide_cast_weights": (_int, [_vp, C.POINTER(WideDesc), _vp, _vp, _vp]),
    "wx_wide_forward": (_int, [_vp, C.POINTER(WideDesc), _vp, _vp, _vp, _vp, _i64, _int, _u64, _u64, C.c_float, _vp,
                               _vp]),
    "wx_wide_backward_layer": (_int, [_vp, C.POINTER(WideDesc), _vp, _i64, _int, _int, _int, _vp, _vp]),
    "wx_wide_adam_layer": (_int, [_vp, C.POINTER(WideDesc), _vp, _vp, _vp, _vp, _i64, _vp, _int, _vp]),
    "wx_wide_adam": (_int, [_vp, C.POINTER(WideDesc), _vp, _vp, _vp, _vp, _i64, _vp, _vp]),
    "wx_wide_reserve_sms": (_int, [_vp, _int]),
    "wx_wide_eval": (_int, [_vp, C.POINTER(WideDesc), _vp, _vp, _vp, _vp, _i64, _vp, _vp]),
    "wx_wide_gemm": (_int, [_vp, _int, _int, _vp, _i64, _vp, _i64, _i64, _i64, _i64, _vp, _vp, _vp]),
}

# developer aids (csrc/wxddp_dev.h): not part of the boundary, bound for scripts/ and the primitives test only
DEV_SIGNATURES = {
    "wxdev_cta_times": (_int, [_vp, _vp]),
    "wxdev_tc_selftest": (_int, [_vp, _vp, _vp, _vp, _int, _vp, _vp, _vp, _vp]),
    "wxdev_tc_mma_bench": (_int, [_vp, _int, _int, _int, _vp, _vp]),
}

_lib = None


def lib_path() -> str:
    return _build.LIB_PATH


def load() -> C.CDLL:
    """dlopen libwxddp.so and attach the prototypes.  Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    path = lib_path()
    if not os.path.exists(path):
        raise WxError(f"{path} is missing: build it with `python __graft_entry__.py` (nvcc, sm_100a). "
                      "wxddp has no CPU or PyTorch fallback.")
    lib = C.CDLL(path)
    for name, (res, args) in list(SIGNATURES.items()) + list(DEV_SIGNATURES.items()):
        fn = getattr(lib, name)      # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc: int) -> None:
    if rc != 0:
        msg = load().wx_last_error().decode("utf-8", "replace")
        raise WxError(f"libwxddp error {rc}: {msg}")


def ptr_array(ptrs) -> C.Array:
    """HOST array of device pointers (wx_col_stats / wx_normalize_pack `cols`, `validity`)."""
    arr = (_vp * len(ptrs))()
    for i, p in enumerate(ptrs):
        arr[i] = p
    return arr
