"""numpy front-end of the C warp oracle (oracle/resample_oracle.c).

TEST INFRASTRUCTURE ONLY -- see oracle/__init__.py.
"""
import ctypes

import numpy as np

from . import build_oracle

_lib = None


def _get():
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(build_oracle.build())
        fp = ctypes.POINTER(ctypes.c_float)
        _lib.r2d_forward.argtypes = [fp, fp, fp] + [ctypes.c_int] * 6
        _lib.r2d_backward.argtypes = [fp, fp, fp, fp, fp] + [ctypes.c_int] * 7
        _lib.gsw_forward.argtypes = [fp, fp, fp] + [ctypes.c_int] * 4
        for f in (_lib.r2d_forward, _lib.r2d_backward, _lib.gsw_forward):
            f.restype = None
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(ctypes.POINTER(ctypes.c_float))


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def resample2d_forward(img, flow):
    """ref: imaginaire/third_party/resample2d/src/resample2d_kernel.cu:15-75"""
    img, flow = _f32(img), _f32(flow)
    b, c, hi, wi = img.shape
    _, _, h, w = flow.shape
    out = np.empty((b, c, h, w), np.float32)
    _get().r2d_forward(_p(img), _p(flow), _p(out), b, c, h, w, hi, wi)
    return out


def resample2d_backward(img, flow, gout, trunc_alpha=True):
    """ref: resample2d_kernel.cu:78-129 (d/d image), :131-203 (d/d flow)"""
    img, flow, gout = _f32(img), _f32(flow), _f32(gout)
    b, c, hi, wi = img.shape
    _, _, h, w = flow.shape
    gimg = np.zeros_like(img)
    gflow = np.empty_like(flow)
    _get().r2d_backward(_p(img), _p(flow), _p(gout), _p(gimg), _p(gflow), b, c, h, w, hi, wi, int(trunc_alpha))
    return gimg, gflow


def gridsample_warp_forward(img, flow):
    """ref: imaginaire/model_utils/fs_vid2vid.py:14-39 resample()"""
    img, flow = _f32(img), _f32(flow)
    b, c, h, w = img.shape
    out = np.empty_like(img)
    _get().gsw_forward(_p(img), _p(flow), _p(out), b, c, h, w)
    return out
