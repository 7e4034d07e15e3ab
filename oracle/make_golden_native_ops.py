"""Generates tests/golden/ref_{resample2d,channelnorm,correlation}_*.npz from the REFERENCE's own CUDA kernel
bodies run on host cores (oracle/ref_cuda/: resample2d_kernel.cu:15-203, channelnorm_kernel.cu:18-96,
correlation_cuda_kernel.cu:16-333 compiled with g++ behind a CUDA-on-host shim).

TEST INFRASTRUCTURE ONLY.  Runs only in the build container (needs /root/reference); the fixtures are
committed.  Every output is stored twice: `*_strict` from the -ffp-contract=off build (the source's arithmetic
as written) and `*_fmad` from the contracted build (what nvcc's default -fmad=true makes of it).

    python -m oracle.make_golden_native_ops
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)

from oracle.make_golden import save, synthetic_flow  # noqa: E402
from oracle.ref_cuda import ref_kernels as R  # noqa: E402


def both(fn, *args):
    out = {}
    for tag, fmad in (('strict', False), ('fmad', True)):
        with R.use(fmad):
            res = fn(*args)
        out[tag] = res if isinstance(res, tuple) else (res,)
    return out


def main():
    assert R.available(), 'needs /root/reference'
    rng = np.random.default_rng(2024)

    # ---- resample2d: SURVEY flow set (integer, +-0.5, far out of range, negative coordinates -> the
    #      trunc-vs-floor quirk of resample2d_kernel.cu:109-110 is exercised by the rows with xf < 0)
    for tag, (b, c, h, w) in [('a', (2, 3, 40, 56)), ('b', (1, 5, 17, 23))]:
        img = rng.standard_normal((b, c, h, w)).astype(np.float32)
        flow = synthetic_flow(41, b, h, w, scale=6.0).numpy()
        flow[:, 0, 5, :] = -3.25          # xf < 0 near the left border: int(xf) != floor(xf)
        flow[:, 1, 6, :] = -9.75
        gout = rng.standard_normal((b, c, h, w)).astype(np.float32)
        fwd = both(R.resample2d_forward, img, flow)
        bwd = both(R.resample2d_backward, img, flow, gout)
        save('ref_resample2d_%s.npz' % tag, img=img, flow=flow, gout=gout,
             out_strict=fwd['strict'][0], out_fmad=fwd['fmad'][0],
             gimg_strict=bwd['strict'][0], gflow_strict=bwd['strict'][1],
             gimg_fmad=bwd['fmad'][0], gflow_fmad=bwd['fmad'][1])

    # ---- channelnorm
    for tag, shape in [('a', (2, 3, 17, 23)), ('b', (3, 64, 5, 7))]:
        x = rng.standard_normal(shape).astype(np.float32)
        fwd = both(R.channelnorm_forward, x)
        gout = rng.standard_normal(fwd['strict'][0].shape).astype(np.float32)
        save('ref_channelnorm_%s.npz' % tag, x=x, gout=gout, out_strict=fwd['strict'][0], out_fmad=fwd['fmad'][0],
             gin_strict=both(R.channelnorm_backward, x, fwd['strict'][0], gout)['strict'][0],
             gin_fmad=both(R.channelnorm_backward, x, fwd['fmad'][0], gout)['fmad'][0])

    # ---- correlation: FlowNetC geometry (pad 20, k 1, d 20, s1 1, s2 2; flownet_c.py:44-46) on a small map with a
    #      channel count that is not a multiple of the 32 lanes, and two denser displacement grids
    for tag, (b, c, h, w), cfg in [('a', (1, 40, 6, 8), (20, 1, 20, 1, 2)),
                                   ('b', (2, 64, 9, 11), (4, 1, 4, 1, 2)),
                                   ('c', (1, 32, 10, 14), (4, 1, 4, 1, 1))]:
        x1 = rng.standard_normal((b, c, h, w)).astype(np.float32)
        x2 = rng.standard_normal((b, c, h, w)).astype(np.float32)
        fwd = both(R.correlation_forward, x1, x2, *cfg)
        gout = rng.standard_normal(fwd['strict'][0].shape).astype(np.float32)
        bwd = both(R.correlation_backward, x1, x2, gout, *cfg)
        save('ref_correlation_%s.npz' % tag, x1=x1, x2=x2, gout=gout, cfg=np.array(cfg),
             out_strict=fwd['strict'][0], out_fmad=fwd['fmad'][0],
             g1_strict=bwd['strict'][0], g2_strict=bwd['strict'][1], g1_fmad=bwd['fmad'][0], g2_fmad=bwd['fmad'][1])


if __name__ == '__main__':
    main()
