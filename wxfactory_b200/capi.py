"""ctypes binding of libwxb200.so (include/wxb200.h).  No torch types cross this boundary.

The product path has no CPU fallback: if the library is missing or no CUDA device is present the calls
raise (``RuntimeError``), they never route through numpy.
"""

import ctypes
import os

import numpy

from . import build as _build

WXB_MAX_SLABS = 8
EQ_EULER3D, EQ_SW = 0, 1
JVP_FD, JVP_RAT = 0, 1

c_double_p = ctypes.POINTER(ctypes.c_double)
c_int32_p = ctypes.POINTER(ctypes.c_int32)


class WxbConfig(ctypes.Structure):
    _fields_ = [
        ("equations", ctypes.c_int32),
        ("n", ctypes.c_int32),
        ("nz", ctypes.c_int32),
        ("nx", ctypes.c_int32),
        ("nslabs", ctypes.c_int32),
        ("slab_ny", ctypes.c_int32 * WXB_MAX_SLABS),
        ("nghost", ctypes.c_int32),
        ("nsend", ctypes.c_int32),
        ("nxform", ctypes.c_int32),
        ("advection_only", ctypes.c_int32),
        ("has_damping", ctypes.c_int32),
        ("has_topo", ctypes.c_int32),
        ("damp_tau0", ctypes.c_double),
    ]


_ARRAY_FIELDS = [
    ("ext_w", c_double_p),
    ("ext_e", c_double_p),
    ("diff", c_double_p),
    ("corr", c_double_p),
    ("hfilt", c_double_p),
    ("sqrtg", c_double_p),
    ("h", c_double_p),
    ("chris", c_double_p),
    ("inv_dzdeta", c_double_p),
    ("damp", c_double_p),
    ("topo", c_double_p),
    ("fm_i", c_double_p),
    ("fm_j", c_double_p),
    ("fm_k", c_double_p),
    ("topo_itf", c_double_p),
    ("nbr", c_int32_p),
    ("xform", c_double_p),
    ("send", c_int32_p),
]


class WxbStaticArrays(ctypes.Structure):
    _fields_ = _ARRAY_FIELDS


WXB_VEC_MAX_ROWS = 64  # include/wxb200.h
WXB_NCCL_ID_BYTES = 128

EXPORTS = [
    "wxb_version",
    "wxb_create",
    "wxb_destroy",
    "wxb_last_error",
    "wxb_handle_fused",
    "wxb_state_size",
    "wxb_send_size",
    "wxb_recv_size",
    "wxb_send_buffer",
    "wxb_recv_buffer",
    "wxb_rhs",
    "wxb_jvp",
    "wxb_traces",
    "wxb_finish",
    "wxb_comm_unique_id",
    "wxb_comm_init",
    "wxb_exchange",
    "wxb_eval_dist",
    "wxb_rhs_host",
    "wxb_jvp_host",
    "wxb_linearize_host",
    "wxb_bind_host",
    "wxb_jvp_bound_host",
    "wxb_timings",
    "wxb_timing_history",
    "wxb_set_expfilter",
    "wxb_expfilter",
    "wxb_vec_lincomb3",
    "wxb_vec_nonfinite",
    "wxb_vec_multidot_scratch",
    "wxb_vec_multidot",
    "wxb_vec_multidot_static",
    "wxb_vec_reduce",
    "wxb_vec_combine",
    "wxb_vec_gs1sync",
]

_lib = None


def library_path():
    return _build.LIB


def load():
    """Load (building first if the sources are newer) and type the C ABI."""
    global _lib
    if _lib is not None:
        return _lib
    path = os.environ.get("WXB_LIB_PATH") or _build.build_if_stale()  # WXB_LIB_PATH: an alternative build (A/B runs)
    lib = ctypes.CDLL(path)
    vp, i32, i64, dbl = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_double
    lib.wxb_version.restype = i32
    lib.wxb_create.argtypes = [ctypes.POINTER(WxbConfig), ctypes.POINTER(WxbStaticArrays), i32, ctypes.POINTER(vp)]
    lib.wxb_destroy.argtypes = [vp]
    lib.wxb_handle_fused.argtypes = [vp]
    lib.wxb_last_error.argtypes = [vp]
    lib.wxb_last_error.restype = ctypes.c_char_p
    for f in ("wxb_state_size", "wxb_send_size", "wxb_recv_size"):
        getattr(lib, f).argtypes = [vp]
        getattr(lib, f).restype = i64
    for f in ("wxb_send_buffer", "wxb_recv_buffer"):
        getattr(lib, f).argtypes = [vp]
        getattr(lib, f).restype = vp
    lib.wxb_rhs.argtypes = [vp, vp, vp, vp]
    lib.wxb_jvp.argtypes = [vp, vp, vp, vp, dbl, dbl, i32, vp, vp]
    lib.wxb_traces.argtypes = [vp, vp, vp, dbl, vp]
    lib.wxb_finish.argtypes = [vp, vp, vp, vp, dbl, dbl, i32, i32, vp, vp]
    lib.wxb_comm_unique_id.argtypes = [vp]
    lib.wxb_comm_init.argtypes = [vp, vp, i32, i32, i32, vp, vp, vp, i32, vp, vp, vp]
    lib.wxb_exchange.argtypes = [vp, vp]
    lib.wxb_eval_dist.argtypes = [vp, vp, vp, vp, dbl, dbl, i32, vp, vp]
    lib.wxb_rhs_host.argtypes = [vp, vp, vp]
    lib.wxb_jvp_host.argtypes = [vp, vp, vp, vp, dbl, dbl, i32, vp]
    lib.wxb_linearize_host.argtypes = [vp, vp, vp]
    lib.wxb_bind_host.argtypes = [vp, vp, vp]
    lib.wxb_jvp_bound_host.argtypes = [vp, vp, dbl, dbl, i32, vp]
    lib.wxb_timings.argtypes = [vp, ctypes.POINTER(ctypes.c_float * 3)]
    lib.wxb_timing_history.argtypes = [vp, i32, vp, vp]
    lib.wxb_timing_history.restype = i32
    lib.wxb_set_expfilter.argtypes = [vp, vp]
    lib.wxb_expfilter.argtypes = [vp, vp, vp, vp]
    lib.wxb_vec_lincomb3.argtypes = [i64, dbl, vp, dbl, vp, dbl, vp, dbl, vp, vp]
    lib.wxb_vec_nonfinite.argtypes = [i64, vp, vp, vp]
    lib.wxb_vec_multidot_scratch.argtypes = [i32, i32]
    lib.wxb_vec_multidot_scratch.restype = i64
    lib.wxb_vec_multidot.argtypes = [i64, i32, vp, i64, i32, vp, i64, vp, vp, vp]
    lib.wxb_vec_multidot_static.argtypes = [i64, i32, vp, i64, i32, vp, i64, vp, vp]
    lib.wxb_vec_reduce.argtypes = [i64, i32, vp, vp, vp, vp, vp]
    lib.wxb_vec_combine.argtypes = [i64, i32, vp, vp, i64, dbl, vp, dbl, vp, vp]
    lib.wxb_vec_gs1sync.argtypes = [i64, i32, vp, vp, vp, i64, dbl, dbl, vp, vp, vp]
    _lib = lib
    return lib


class WxbError(RuntimeError):
    pass


def check(lib, handle, code, what):
    if code != 0:
        msg = lib.wxb_last_error(handle)
        msg = msg.decode() if msg else ""
        exc = ValueError if code == -1 else WxbError
        raise exc(f"{what} failed ({code}): {msg}")


def as_f64(a):
    return numpy.ascontiguousarray(a, dtype=numpy.float64)


def ptr(a, typ=c_double_p):
    if a is None:
        return typ()
    return a.ctypes.data_as(typ)
