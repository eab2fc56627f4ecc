"""ctypes binding of ``libmsgat_sm100.so`` (C ABI: ``include/msgat_sm100.h``).

The library is hand-written CUDA for sm_100a and is the ONLY implementation of the hot
path: there is no CPU or PyTorch fallback.  ``get()`` raises if the shared object has not
been built (``python -c "import __graft_entry__ as g; g.build()"``) or if no compute
capability 10.x device is present.
"""
from __future__ import annotations

import ctypes as C
import os

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.path.join(HERE, "csrc", "libmsgat_sm100.so")

MAX_CONV = 4
MAX_GROUP = 8
F_LN, F_RES, F_RELU, F_CMIX, F_SMIX, F_DOUT_MASKED, F_X_IS_RELU = 1, 2, 4, 8, 16, 32, 64
PREC_FP32, PREC_TF32 = 0, 1

_fp = C.c_void_p


class MeamDesc(C.Structure):
    _fields_ = [("B", C.c_int32), ("C", C.c_int32), ("N", C.c_int32),
                ("Oc", C.c_int32), ("Ot", C.c_int32), ("Os", C.c_int32),
                ("flags", C.c_int32), ("n_conv", C.c_int32), ("dil", C.c_int32 * MAX_CONV),
                ("precision", C.c_int32), ("M", C.c_int32),
                ("x_sb", C.c_int64), ("x_sc", C.c_int64), ("x_sm", C.c_int64)]


_PARAM_FIELDS = ["ln_w", "ln_b", "res_w", "res_b", "ca_Wc", "ca_alpha", "ca_w", "ca_b",
                 "ta_Wt1", "ta_Wt2", "ta_alpha"]


class MeamParams(C.Structure):
    _fields_ = ([(n, _fp) for n in _PARAM_FIELDS] + [("conv_w", _fp * MAX_CONV), ("conv_b", _fp * MAX_CONV)]
                + [("sa_W", _fp), ("sa_Wg", _fp), ("sa_alpha", _fp)])


class MeamGrads(C.Structure):
    _fields_ = MeamParams._fields_


class Adj(C.Structure):
    _fields_ = [("N", C.c_int32), ("nnz", C.c_int32), ("row_ptr", _fp), ("col_idx", _fp), ("val", _fp),
                ("csc_ptr", _fp), ("csc_row", _fp), ("csc_eid", _fp)]


class HeadDesc(C.Structure):
    _fields_ = [("B", C.c_int32), ("C", C.c_int32), ("N", C.c_int32), ("Tout", C.c_int32),
                ("accumulate", C.c_int32), ("x_is_relu", C.c_int32), ("gate_sb", C.c_int64)]


class HeadsDesc(C.Structure):
    _fields_ = [("B", C.c_int32), ("C", C.c_int32), ("N", C.c_int32), ("Tout", C.c_int32), ("M", C.c_int32),
                ("x_is_relu", C.c_int32), ("x_sm", C.c_int64), ("gate_sb", C.c_int64), ("gate_sm", C.c_int64)]


EXPORTS = ["msgat_version", "msgat_last_error", "msgat_device_check", "msgat_launch_count", "msgat_profile",
           "msgat_profile_report", "msgat_adj_count", "msgat_adj_fill",
           "msgat_meam_sizes", "msgat_meam_forward", "msgat_meam_backward", "msgat_head_forward",
           "msgat_head_scratch_floats", "msgat_head_backward", "msgat_tc_probe",
           "msgat_heads_forward", "msgat_heads_scratch_floats", "msgat_heads_backward",
           "msgat_loss_scratch_bytes", "msgat_loss_forward", "msgat_loss_backward", "msgat_adam_step",
           "msgat_loss_static_max_n", "msgat_loss_forward_static", "msgat_window_gather"]

_LIB = None
_DEBUG_BACKEND = False  # set only by tests/emu (kernel-logic debugging on the GPU-less build box)


def _declare(lib):
    lib.msgat_version.restype = C.c_char_p
    lib.msgat_last_error.restype = C.c_char_p
    lib.msgat_device_check.restype = C.c_int
    lib.msgat_head_scratch_floats.restype = C.c_int64
    lib.msgat_heads_scratch_floats.restype = C.c_int64
    lib.msgat_launch_count.restype = C.c_longlong
    lib.msgat_loss_scratch_bytes.restype = C.c_int64
    lib.msgat_loss_forward.argtypes = [_fp, _fp, C.c_int64, C.c_float, C.c_float, _fp, _fp, _fp, _fp, _fp]
    lib.msgat_loss_backward.argtypes = [_fp, _fp, C.c_int64, _fp, _fp]
    lib.msgat_loss_static_max_n.restype = C.c_int64
    lib.msgat_window_gather.argtypes = [_fp, _fp, _fp, C.c_int, C.c_int, C.c_int, C.c_int64, C.POINTER(C.c_int32), C.c_int,
                                        C.c_int, C.c_int, _fp, _fp, _fp, _fp, _fp]
    lib.msgat_loss_forward_static.argtypes = [_fp, _fp, C.c_int64, C.c_float, C.c_float, _fp, _fp, _fp, _fp]
    lib.msgat_adam_step.argtypes = [_fp, _fp, _fp, _fp, C.c_int64, C.c_float, C.c_float, C.c_float, C.c_float,
                                    C.c_float, C.c_float, _fp, _fp, _fp]
    for name in EXPORTS:
        getattr(lib, name)  # AttributeError if the .so is stale
    return lib


def load_unchecked(path: str = SO_PATH):
    """dlopen the library and bind every export, without touching a device (used by the
    CPU-side symbol test)."""
    if not os.path.exists(path):
        raise RuntimeError(f"{path} is missing: build it with __graft_entry__.build(); there is no fallback path")
    return _declare(C.CDLL(path))


def get():
    global _LIB
    if _LIB is None:
        if not torch.cuda.is_available():
            raise RuntimeError("msgat_b200 needs a CUDA device of compute capability 10.x (B200); "
                               "there is no CPU fallback")
        lib = load_unchecked()
        if lib.msgat_device_check() != 0:
            raise RuntimeError("msgat_b200: " + lib.msgat_last_error().decode())
        _LIB = lib
    return _LIB


def check(rc: int, what: str):
    if rc != 0:
        raise RuntimeError(f"{what} failed ({rc}): {get().msgat_last_error().decode()}")


def ptr(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def stream_of(t: torch.Tensor):
    if t.is_cuda:
        return C.c_void_p(torch.cuda.current_stream(t.device).cuda_stream)
    if not _DEBUG_BACKEND:
        raise RuntimeError("msgat_b200 operators take CUDA tensors only")
    return C.c_void_p(0)


def launch_count() -> int:
    return int(get().msgat_launch_count())


def profile(enable: bool):
    get().msgat_profile(1 if enable else 0)


def profile_report() -> dict:
    """{kernel: (launches, total_ms)} of the launches recorded since profile(True)."""
    buf = C.create_string_buffer(1 << 16)
    check(get().msgat_profile_report(buf, len(buf)), "msgat_profile_report")
    out = {}
    for line in buf.value.decode().splitlines():
        name, cnt, ms = line.rsplit(" ", 2)
        out[name] = (int(cnt), float(ms))
    return out
