"""ctypes binding of libbt_b200.so (the C ABI declared in include/bt_b200.h).

There is no fallback: if the shared library is missing or a call fails the
Python layer raises -- the CUDA path is the only product path.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "csrc", "libbt_b200.so")

BT_F32, BT_BF16 = 0, 1
PATH_AUTO, PATH_SIMT_F32, PATH_TC_BF16 = 0, 1, 2
KIND_EPS_W, KIND_EPS_B, KIND_SIGN_IN, KIND_SIGN_OUT = 0, 1, 2, 3
FLAG_PACKED_VALID = 1
FLAG_TRANSPOSE_BIAS = 2
FLAG_BIAS_GRAD_X_SIDE = 4
ABI_VERSION = 5

_i64x3 = C.c_int64 * 3


class Geom(C.Structure):
    _fields_ = [
        ("rank", C.c_int32), ("groups", C.c_int32), ("n_samples", C.c_int32),
        ("x_dtype", C.c_int32), ("out_dtype", C.c_int32), ("reserved", C.c_int32),
        ("batch", C.c_int64), ("c_in", C.c_int64), ("c_out", C.c_int64),
        ("in_size", _i64x3), ("out_size", _i64x3), ("k_size", _i64x3),
        ("stride", _i64x3), ("pad", _i64x3), ("dil", _i64x3),
        ("x_sample_stride", C.c_int64), ("x_batch_stride", C.c_int64), ("x_chan_stride", C.c_int64),
        ("x_sp_stride", _i64x3),
        ("out_sample_stride", C.c_int64), ("out_batch_stride", C.c_int64), ("out_chan_stride", C.c_int64),
        ("out_sp_stride", _i64x3),
    ]


class LayerParams(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in (
        "mu_w", "rho_w", "mu_b", "rho_b", "eps_w", "eps_b", "sign_in", "sign_out",
        "prior_mu_w_t", "prior_sigma_w_t", "prior_mu_b_t", "prior_sigma_b_t")] + [
        ("prior_mu", C.c_float), ("prior_sigma", C.c_float),
        ("seed", C.c_uint64), ("sample_id", C.c_uint32), ("layer_id", C.c_uint32),
        ("out_scale", C.c_void_p), ("out_shift", C.c_void_p), ("out_relu", C.c_int32), ("flags", C.c_int32),
        ("residual", C.c_void_p), ("workspace_bytes", C.c_int64)]


class LstmLayer(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ("mu_w", "rho_w", "mu_b", "rho_b", "eps_w", "eps_b", "sign_in", "sign_out")] + [
        ("seed", C.c_uint64), ("sample_id", C.c_uint32), ("layer_id", C.c_uint32)]


class LstmStep(C.Structure):
    _fields_ = [("batch", C.c_int64), ("in_features", C.c_int32), ("hidden", C.c_int32), ("flipout", C.c_int32),
                ("reserved", C.c_int32), ("ih", LstmLayer), ("hh", LstmLayer),
                ("x", C.c_void_p), ("x_row_stride", C.c_int64), ("h_prev", C.c_void_p), ("h_row_stride", C.c_int64),
                ("c_prev", C.c_void_p), ("c_row_stride", C.c_int64), ("h_out", C.c_void_p), ("h_out_row_stride", C.c_int64),
                ("c_out", C.c_void_p), ("c_out_row_stride", C.c_int64), ("gates_out", C.c_void_p)]


class LayerGrads(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ("grad_x", "grad_mu_w", "grad_rho_w", "grad_mu_b", "grad_rho_b")]


_FWD = [C.POINTER(Geom), C.POINTER(LayerParams), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
_BWD = [C.POINTER(Geom), C.POINTER(LayerParams), C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(LayerGrads),
        C.c_void_p, C.c_int, C.c_void_p]

#: every symbol include/bt_b200.h declares -> (restype, argtypes)
SIGNATURES = {
    "bt_abi_version": (C.c_int, []),
    "bt_last_error": (C.c_char_p, []),
    "bt_shutdown": (C.c_int, []),
    "bt_launch_count": (C.c_int64, []),
    "bt_workspace_bytes": (C.c_int64, [C.c_char_p, C.POINTER(Geom), C.c_int]),
    "bt_select_path": (C.c_int, [C.POINTER(Geom), C.POINTER(LayerParams), C.c_int, C.c_int]),
    "bt_kl_normal_multi": (C.c_int, [C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), C.POINTER(C.c_int64),
                                     C.c_int, C.c_float, C.c_float, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p),
                                     C.c_void_p, C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p),
                                     C.c_void_p, C.c_void_p, C.c_void_p]),
    "bt_kl_normal_multi_static": (C.c_int, [C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), C.POINTER(C.c_int64),
                                            C.c_int, C.c_float, C.c_float, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p),
                                            C.c_void_p, C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p),
                                            C.c_void_p, C.c_void_p]),
    "bt_kl_laplace_multi": (C.c_int, [C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), C.POINTER(C.c_int64),
                                      C.c_int, C.c_void_p, C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p),
                                      C.c_void_p, C.c_void_p, C.c_void_p]),
    "bt_linear_reparam_fwd": (C.c_int, _FWD),
    "bt_conv_reparam_fwd": (C.c_int, _FWD),
    "bt_linear_flipout_fwd": (C.c_int, _FWD),
    "bt_conv2d_flipout_fwd": (C.c_int, _FWD),
    "bt_linear_reparam_bwd": (C.c_int, _BWD),
    "bt_conv_reparam_bwd": (C.c_int, _BWD),
    "bt_linear_flipout_bwd": (C.c_int, _BWD),
    "bt_conv2d_flipout_bwd": (C.c_int, _BWD),
    "bt_lstm_step_fwd": (C.c_int, [C.POINTER(LstmStep), C.c_void_p]),
    "bt_lstm_step_bwd_gates": (C.c_int, [C.c_int64, C.c_int, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64,
                                         C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]),
    "bt_mc_accumulate": (C.c_int, [C.c_void_p, C.c_int, C.c_int64, C.c_int64, C.c_int64, C.c_int,
                                   C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "bt_mc_finalize": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64,
                                 C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "bt_philox_normal_fill": (C.c_int, [C.c_void_p, C.c_int64, C.c_uint64, C.c_uint32, C.c_uint32, C.c_int, C.c_void_p]),
    "bt_philox_sign_fill": (C.c_int, [C.c_void_p, C.c_int64, C.c_uint64, C.c_uint32, C.c_uint32, C.c_int, C.c_void_p]),
    "bt_moped_rho": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_float, C.c_void_p]),
    "bt_add_relu": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_void_p]),
    "bt_maxpool2d_nhwc": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_int64,
                                    C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]),
    "bt_debug_tc_profile": (C.c_int, [C.c_void_p, C.c_int]),
}

_lib = None


def lib() -> C.CDLL:
    """Load (once) and return the shared library; raises if it was not built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python bayesian-torch_b200/csrc/build.py` "
                "(or __graft_entry__.build()). There is no CPU / eager fallback.")
        handle = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)
            fn.restype = res
            fn.argtypes = args
        if handle.bt_abi_version() != ABI_VERSION:
            raise RuntimeError("libbt_b200.so ABI version mismatch")
        _lib = handle
    return _lib


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = lib().bt_last_error().decode("utf-8", "replace")
        raise RuntimeError(f"bt_b200 {what} failed (status {rc}): {msg}")
