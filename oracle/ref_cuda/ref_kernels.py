"""numpy front-end of oracle/_ref/libref_cuda_kernels.so (the reference's own kernel bodies run on host cores).

TEST INFRASTRUCTURE ONLY -- see oracle/ref_cuda/build_ref.py.  Same call signatures as
oracle/warp_oracle.py / oracle/flownet_ops_oracle.py, so a test can hold the restatements to it.
"""
import ctypes

import numpy as np

from . import build_ref

_libs = {}
_FMAD = False     # module-level switch: use(fmad=True) selects the nvcc-like (contracted) build


def available():
    return build_ref.build() is not None


class use:
    """Context manager: ``with use(fmad=True): ...`` runs the calls inside on the -mfma build."""

    def __init__(self, fmad):
        self.fmad = fmad

    def __enter__(self):
        global _FMAD
        self.prev, _FMAD = _FMAD, self.fmad

    def __exit__(self, *exc):
        global _FMAD
        _FMAD = self.prev


def _get():
    if _FMAD not in _libs:
        path = build_ref.build()
        if path is None:
            raise RuntimeError('oracle/_ref/libref_cuda_kernels.so is not built and /root/reference is absent')
        _lib = ctypes.CDLL(build_ref.LIB_FMAD if _FMAD else path)
        fp, ip, ci = ctypes.POINTER(ctypes.c_float), ctypes.POINTER(ctypes.c_int), ctypes.c_int
        _lib.ref_resample2d_forward.argtypes = [fp] * 3 + [ci] * 7
        _lib.ref_resample2d_backward.argtypes = [fp] * 5 + [ci] * 7
        _lib.ref_channelnorm_forward.argtypes = [fp] * 2 + [ci] * 5
        _lib.ref_channelnorm_backward.argtypes = [fp] * 4 + [ci] * 5
        _lib.ref_correlation_output_size.argtypes = [ci] * 8 + [ip] * 3
        _lib.ref_correlation_forward.argtypes = [fp] * 3 + [ci] * 9
        _lib.ref_correlation_backward.argtypes = [fp] * 5 + [ci] * 9
        for name in ('ref_resample2d_forward', 'ref_resample2d_backward', 'ref_channelnorm_forward',
                     'ref_channelnorm_backward', 'ref_correlation_output_size', 'ref_correlation_forward',
                     'ref_correlation_backward'):
            getattr(_lib, name).restype = None
        _libs[_FMAD] = _lib
    return _libs[_FMAD]


def _p(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_float))


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def resample2d_forward(img, flow):
    img, flow = _f32(img), _f32(flow)
    b, c, hi, wi = img.shape
    _, _, h, w = flow.shape
    out = np.zeros((b, c, h, w), np.float32)        # resample2d.py:21 zero-fills
    _get().ref_resample2d_forward(_p(img), _p(flow), _p(out), b, c, h, w, hi, wi, 1)
    return out


def resample2d_backward(img, flow, gout):
    img, flow, gout = _f32(img), _f32(flow), _f32(gout)
    b, c, hi, wi = img.shape
    _, _, h, w = flow.shape
    gimg, gflow = np.zeros_like(img), np.zeros_like(flow)   # resample2d.py:34-35
    _get().ref_resample2d_backward(_p(img), _p(flow), _p(gout), _p(gimg), _p(gflow), b, c, h, w, hi, wi, 1)
    return gimg, gflow


def channelnorm_forward(x):
    x = _f32(x)
    b, c, h, w = x.shape
    out = np.zeros((b, 1, h, w), np.float32)
    _get().ref_channelnorm_forward(_p(x), _p(out), b, c, h, w, 2)
    return out


def channelnorm_backward(x, out, gout):
    x, out, gout = _f32(x), _f32(out), _f32(gout)
    b, c, h, w = x.shape
    gin = np.zeros_like(x)
    _get().ref_channelnorm_backward(_p(x), _p(out), _p(gout), _p(gin), b, c, h, w, 2)
    return gin


def correlation_output_size(c, h, w, pad, ksize, max_disp, stride1, stride2):
    oc, oh, ow = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
    _get().ref_correlation_output_size(c, h, w, pad, ksize, max_disp, stride1, stride2, ctypes.byref(oc),
                                       ctypes.byref(oh), ctypes.byref(ow))
    return oc.value, oh.value, ow.value


def correlation_forward(x1, x2, pad, ksize, max_disp, stride1, stride2):
    x1, x2 = _f32(x1), _f32(x2)
    b, c, h, w = x1.shape
    oc, oh, ow = correlation_output_size(c, h, w, pad, ksize, max_disp, stride1, stride2)
    out = np.empty((b, oc, oh, ow), np.float32)
    _get().ref_correlation_forward(_p(x1), _p(x2), _p(out), b, c, h, w, pad, ksize, max_disp, stride1, stride2)
    return out


def correlation_backward(x1, x2, gout, pad, ksize, max_disp, stride1, stride2):
    x1, x2, gout = _f32(x1), _f32(x2), _f32(gout)
    b, c, h, w = x1.shape
    g1, g2 = np.empty_like(x1), np.empty_like(x2)
    _get().ref_correlation_backward(_p(x1), _p(x2), _p(gout), _p(g1), _p(g2), b, c, h, w, pad, ksize, max_disp,
                                    stride1, stride2)
    return g1, g2
