"""ctypes loader for libfetode.so (the C ABI declared in include/fetode.h).

There is no CPU fallback: if the shared library is missing this module raises at import
of any compute entry point, and every non-zero return code becomes a RuntimeError (usage
errors that the reference reports as AssertionError/ValueError are raised as such by the
module layer *before* the call).
"""
from __future__ import annotations

import ctypes as C
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libfetode.so")

BRANCH_NONE, BRANCH_LOGISTIC, BRANCH_FERRO2D = 0, 1, 2
TABLEAU = {"euler": 0, "midpoint": 1, "rk4_classic": 2, "rk4_38": 3}
FERRO_FULL, FERRO_TWO_D, FERRO_HARDGATE = 0, 1, 2

c_float_p = C.POINTER(C.c_float)


class KanLayer(C.Structure):
    _fields_ = [
        ("in_features", C.c_int32), ("out_features", C.c_int32),
        ("grid_size", C.c_int32), ("spline_order", C.c_int32),
        ("num_basis", C.c_int32), ("branch", C.c_int32),
        ("grid", C.c_void_p), ("base_weight", C.c_void_p), ("spline_weight", C.c_void_p),
        ("spline_scaler", C.c_void_p), ("branch_weight", C.c_void_p),
        ("branch_scaler", C.c_void_p), ("branch_bias", C.c_void_p),
        ("bp", C.c_void_p * 5),
        ("scale_logistic", C.c_float), ("gate_slope", C.c_float), ("alpha", C.c_float),
    ]


class KanLayerGrads(C.Structure):
    _fields_ = [
        ("base_weight", C.c_void_p), ("spline_weight", C.c_void_p), ("spline_scaler", C.c_void_p),
        ("branch_weight", C.c_void_p), ("branch_scaler", C.c_void_p), ("branch_bias", C.c_void_p),
        ("bp", C.c_void_p * 5),
    ]


class SmallNet(C.Structure):
    _fields_ = [
        ("n_layers", C.c_int32), ("widths", C.c_int32 * 3),
        ("grid_size", C.c_int32), ("spline_order", C.c_int32),
        ("num_basis", C.c_int32), ("branch", C.c_int32),
    ]


class Ferro(C.Structure):
    _fields_ = [
        ("variant", C.c_int32), ("in_dim", C.c_int32), ("out_dim", C.c_int32), ("num_basis", C.c_int32),
        ("k", C.c_void_p), ("Ec", C.c_void_p), ("Ps", C.c_void_p), ("bias", C.c_void_p), ("coef", C.c_void_p),
        ("gate_slope", C.c_float), ("alpha", C.c_float), ("apply_sigmoid", C.c_int32),
        ("pre_bound", C.c_float), ("post_op", C.c_int32),
    ]


class FerroGrads(C.Structure):
    _fields_ = [("k", C.c_void_p), ("Ec", C.c_void_p), ("Ps", C.c_void_p), ("bias", C.c_void_p),
                ("coef", C.c_void_p)]


_lib = None


def lib() -> C.CDLL:
    """Load libfetode.so or fail loudly (no fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} not found: build it with `python fet-ode_b200/build.py` "
                "(fetode has no CPU or PyTorch fallback)")
        _lib = C.CDLL(LIB_PATH)
        _lib.fetode_strerror.restype = C.c_char_p
        _lib.fetode_strerror.argtypes = [C.c_int]
        for name in ("fetode_kanlinear_wcat_floats", "fetode_kanlinear_bwd_ws_floats",
                     "fetode_smallnet_packed_floats", "fetode_smallnet_state_floats",
                     "fetode_smallnet_ckpt_floats", "fetode_smallnet_gfold_floats",
                     "fetode_smallnet_wide_ckpt_floats", "fetode_smallnet_attempt_ckpt_floats"):
            if hasattr(_lib, name):
                getattr(_lib, name).restype = C.c_int64
    return _lib


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = lib().fetode_strerror(int(rc)).decode()
        raise RuntimeError(f"{what or 'fetode'}: {msg} (rc={rc})")


def ptr(t: torch.Tensor | None) -> C.c_void_p:
    """Device pointer of a contiguous fp32 CUDA tensor (None -> NULL)."""
    if t is None:
        return C.c_void_p(0)
    if not t.is_cuda:
        raise RuntimeError("fetode: tensors must live on a CUDA device (there is no CPU path)")
    if not t.is_contiguous():
        raise RuntimeError("fetode: tensor must be contiguous")
    return C.c_void_p(t.data_ptr())


def stream() -> C.c_void_p:
    """cudaStream_t of torch's current stream on the current device (raw handle, no Stream object built)."""
    return C.c_void_p(torch._C._cuda_getCurrentRawStream(torch._C._cuda_getDevice()))


def f32c(t: torch.Tensor) -> torch.Tensor:
    """Contiguous fp32 view/copy (plumbing only)."""
    if t.dtype != torch.float32:
        t = t.float()
    return t.contiguous()


class RkAttempt(C.Structure):
    """fetode_rk_attempt_t (include/fetode.h)."""
    _fields_ = [("n_stages", C.c_int32), ("first_eval", C.c_int32), ("y1_is_last_input", C.c_int32), ("_pad", C.c_int32),
                ("a", (C.c_float * 7) * 7), ("e", C.c_float * 7), ("dt", C.c_float), ("rtol", C.c_float),
                ("atol", C.c_float)]


class Conv2d(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("Cin", "H", "W", "kH", "kW", "stride", "padding", "Hout", "Wout")]
