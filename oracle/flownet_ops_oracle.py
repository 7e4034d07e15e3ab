"""numpy front-end of the C oracle for channelnorm / correlation (oracle/flownet_ops_oracle.c).

TEST INFRASTRUCTURE ONLY -- see oracle/__init__.py.
"""
import ctypes

import numpy as np

from . import build_oracle

_lib = None


def _get():
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(build_oracle.build_flownet())
        fp, ip, ci = ctypes.POINTER(ctypes.c_float), ctypes.POINTER(ctypes.c_int), ctypes.c_int
        _lib.cn_forward.argtypes = [fp, fp] + [ci] * 4
        _lib.cn_backward.argtypes = [fp, fp, fp, fp] + [ci] * 4
        _lib.corr_output_size.argtypes = [ci] * 8 + [ip] * 3
        _lib.corr_forward.argtypes = [fp, fp, fp] + [ci] * 9
        _lib.corr_backward.argtypes = [fp] * 5 + [ci] * 9
        for f in (_lib.cn_forward, _lib.cn_backward, _lib.corr_output_size, _lib.corr_forward, _lib.corr_backward):
            f.restype = None
    return _lib


def _p(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_float))


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def channelnorm_forward(x):
    """ref: channelnorm_kernel.cu:18-61"""
    x = _f32(x)
    b, c, h, w = x.shape
    out = np.empty((b, 1, h, w), np.float32)
    _get().cn_forward(_p(x), _p(out), b, c, h, w)
    return out


def channelnorm_backward(x, out, gout):
    """ref: channelnorm_kernel.cu:64-96"""
    x, out, gout = _f32(x), _f32(out), _f32(gout)
    b, c, h, w = x.shape
    gin = np.empty_like(x)
    _get().cn_backward(_p(x), _p(out), _p(gout), _p(gin), b, c, h, w)
    return gin


def correlation_output_size(c, h, w, pad, ksize, max_disp, stride1, stride2):
    """ref: correlation_cuda.cc:25-35"""
    oc, oh, ow = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
    _get().corr_output_size(c, h, w, pad, ksize, max_disp, stride1, stride2, ctypes.byref(oc), ctypes.byref(oh), ctypes.byref(ow))
    return oc.value, oh.value, ow.value


def correlation_forward(x1, x2, pad, ksize, max_disp, stride1, stride2):
    """ref: correlation_cuda_kernel.cu:72-149"""
    x1, x2 = _f32(x1), _f32(x2)
    b, c, h, w = x1.shape
    oc, oh, ow = correlation_output_size(c, h, w, pad, ksize, max_disp, stride1, stride2)
    out = np.empty((b, oc, oh, ow), np.float32)
    _get().corr_forward(_p(x1), _p(x2), _p(out), b, c, h, w, pad, ksize, max_disp, stride1, stride2)
    return out


def correlation_backward(x1, x2, gout, pad, ksize, max_disp, stride1, stride2):
    """ref: correlation_cuda_kernel.cu:152-336 (stride1 == 1)"""
    assert stride1 == 1
    x1, x2, gout = _f32(x1), _f32(x2), _f32(gout)
    b, c, h, w = x1.shape
    g1, g2 = np.empty_like(x1), np.empty_like(x2)
    _get().corr_backward(_p(x1), _p(x2), _p(gout), _p(g1), _p(g2), b, c, h, w, pad, ksize, max_disp, stride1, stride2)
    return g1, g2
