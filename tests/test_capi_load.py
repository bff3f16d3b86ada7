"""The C-ABI library builds, loads and exports every symbol include/wxb200.h declares (no compute, no GPU)."""

import ctypes
import os
import re

from wxfactory_b200 import build, capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, "include", "wxb200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(wxb_[a-z0-9_]+)\s*\(", text)))


def test_library_builds_and_exports_every_declared_symbol():
    path = build.build()
    assert os.path.exists(path)
    lib = ctypes.CDLL(path)
    names = _declared()
    assert set(names) == set(capi.EXPORTS), (names, capi.EXPORTS)
    for name in names:
        assert hasattr(lib, name), name


def test_version_and_error_convention_without_gpu():
    lib = capi.load()
    assert lib.wxb_version() == 1
    # null arguments are rejected with WXB_EINVAL and a message, never a crash
    assert lib.wxb_create(None, None, 0, None) == -1
    assert b"null" in lib.wxb_last_error(None)
    assert lib.wxb_rhs(None, None, None, None) == -1
    assert lib.wxb_destroy(None) == 0


def test_create_fails_loudly_without_cuda_device():
    import numpy
    import torch

    if torch.cuda.is_available():
        return  # covered by the gpu tests
    cfg = capi.WxbConfig()
    cfg.equations, cfg.n, cfg.nz, cfg.nx, cfg.nslabs = 0, 3, 2, 2, 1
    cfg.slab_ny[0] = 2
    arrs = capi.WxbStaticArrays()
    h = ctypes.c_void_p()
    code = lib_create(cfg, arrs, h)
    assert code == -2  # WXB_ECUDA: no device, and no CPU fallback
    assert b"no CPU fallback" in capi.load().wxb_last_error(None)


def lib_create(cfg, arrs, h):
    return capi.load().wxb_create(ctypes.byref(cfg), ctypes.byref(arrs), 0, ctypes.byref(h))
