"""resample2d / channelnorm / correlation against the REFERENCE's own CUDA kernel bodies.

tests/golden/ref_*.npz were produced by compiling the reference's device code for host cores
(oracle/ref_cuda/, oracle/make_golden_native_ops.py) in two builds: `strict` (-ffp-contract=off) and `fmad`
(products feeding an accumulation fused, as nvcc's default -fmad=true does).  What is held, and to which build:

  * resample2d restatement (oracle/resample_oracle.c) and kernels (csrc/warp.cu): bit-exact to `strict`
    (forward, flow gradient; the image gradient's atomics have no fixed order on a GPU -> 1e-5), within
    north_star's 1e-5 relative error of `fmad`;
  * channelnorm / correlation restatement (oracle/flownet_ops_oracle.c) and kernels (csrc/flownet_ops.cu):
    bit-exact to `fmad`, within 1e-6 of `strict`.

The CPU half pins the restatements (here and, when oracle/_ref is present, on fresh random inputs); the GPU half
(-m gpu) calls the CUDA kernels through the C ABI wrappers.
"""
import numpy as np
import pytest
import torch

from oracle import flownet_ops_oracle as fo
from oracle import warp_oracle as wo
from tests import helpers as H

DEV = 'cuda'
R2D, CN, CORR = ['a', 'b'], ['a', 'b'], ['a', 'b', 'c']


def _close(a, b, tol):
    return np.abs(a - b).max() <= tol * max(np.abs(b).max(), 1e-30)


# ------------------------------------------------------------------ CPU: the restatements

@pytest.mark.parametrize('tag', R2D)
def test_resample2d_restatement_equals_reference_kernels(tag):
    d = H.load_golden('ref_resample2d_%s.npz' % tag)
    out = wo.resample2d_forward(d['img'], d['flow'])
    assert np.array_equal(out, d['out_strict'])
    assert _close(out, d['out_fmad'], 1e-5)
    gi, gf = wo.resample2d_backward(d['img'], d['flow'], d['gout'], trunc_alpha=True)
    assert np.array_equal(gi, d['gimg_strict']) and np.array_equal(gf, d['gflow_strict'])   # sequential scatter: same order
    assert _close(gi, d['gimg_fmad'], 1e-5) and _close(gf, d['gflow_fmad'], 1e-5)
    # the floor-consistent variant differs from the reference exactly where a sampling coordinate is negative
    gi_floor, _ = wo.resample2d_backward(d['img'], d['flow'], d['gout'], trunc_alpha=False)
    assert not np.array_equal(gi_floor, d['gimg_strict'])


@pytest.mark.parametrize('tag', CN)
def test_channelnorm_restatement_equals_reference_kernels(tag):
    d = H.load_golden('ref_channelnorm_%s.npz' % tag)
    out = fo.channelnorm_forward(d['x'])
    assert np.array_equal(out, d['out_fmad']) and _close(out, d['out_strict'], 1e-6)
    gin = fo.channelnorm_backward(d['x'], out, d['gout'])
    assert np.array_equal(gin, d['gin_fmad']) and _close(gin, d['gin_strict'], 1e-6)


@pytest.mark.parametrize('tag', CORR)
def test_correlation_restatement_equals_reference_kernels(tag):
    d = H.load_golden('ref_correlation_%s.npz' % tag)
    cfg = [int(v) for v in d['cfg']]
    out = fo.correlation_forward(d['x1'], d['x2'], *cfg)
    assert np.array_equal(out, d['out_fmad']) and _close(out, d['out_strict'], 1e-6)
    g1, g2 = fo.correlation_backward(d['x1'], d['x2'], d['gout'], *cfg)
    assert np.array_equal(g1, d['g1_fmad']) and np.array_equal(g2, d['g2_fmad'])
    assert _close(g1, d['g1_strict'], 1e-6) and _close(g2, d['g2_strict'], 1e-6)


def test_restatements_equal_reference_kernels_on_fresh_inputs():
    """Where the host build of the reference kernels is present (build container; it also travels to the GPU box),
    compare on inputs that are in no fixture."""
    from oracle.ref_cuda import ref_kernels as R
    if not R.available():
        pytest.skip('oracle/_ref is not built on this machine')
    rng = np.random.default_rng(7)
    for b, c, h, w in [(1, 2, 9, 13), (2, 4, 24, 31)]:
        img = rng.standard_normal((b, c, h, w)).astype(np.float32)
        flow = (rng.standard_normal((b, 2, h, w)) * 5).astype(np.float32)
        gout = rng.standard_normal((b, c, h, w)).astype(np.float32)
        assert np.array_equal(wo.resample2d_forward(img, flow), R.resample2d_forward(img, flow))
        for a, r in zip(wo.resample2d_backward(img, flow, gout, True), R.resample2d_backward(img, flow, gout)):
            assert np.array_equal(a, r)
    with R.use(fmad=True):
        x = rng.standard_normal((2, 5, 7, 9)).astype(np.float32)
        assert np.array_equal(fo.channelnorm_forward(x), R.channelnorm_forward(x))
        x1 = rng.standard_normal((1, 33, 7, 9)).astype(np.float32)
        x2 = rng.standard_normal((1, 33, 7, 9)).astype(np.float32)
        cfg = (3, 1, 3, 1, 1)
        out = R.correlation_forward(x1, x2, *cfg)
        assert np.array_equal(fo.correlation_forward(x1, x2, *cfg), out)
        gout = rng.standard_normal(out.shape).astype(np.float32)
        for a, r in zip(fo.correlation_backward(x1, x2, gout, *cfg), R.correlation_backward(x1, x2, gout, *cfg)):
            assert np.array_equal(a, r)


# ------------------------------------------------------------------ GPU: the kernels, through the C ABI

@pytest.mark.gpu
@pytest.mark.parametrize('tag', R2D)
def test_resample2d_kernels_equal_reference_kernels(tag):
    from imaginaire_b200 import ops
    d = H.load_golden('ref_resample2d_%s.npz' % tag)
    img, flow, gout = (torch.from_numpy(d[k]).to(DEV) for k in ('img', 'flow', 'gout'))
    out = ops.resample2d_fwd(img, flow).cpu().numpy()
    assert np.array_equal(out, d['out_strict']) and _close(out, d['out_fmad'], 1e-5)
    gi, gf = ops.resample2d_bwd(img, flow, gout, trunc_alpha=True)
    assert np.array_equal(gf.cpu().numpy(), d['gflow_strict'])
    assert _close(gi.cpu().numpy(), d['gimg_strict'], 1e-5) and _close(gi.cpu().numpy(), d['gimg_fmad'], 1e-5)


@pytest.mark.gpu
@pytest.mark.parametrize('tag', CN)
def test_channelnorm_kernels_equal_reference_kernels(tag):
    from imaginaire_b200 import ops
    d = H.load_golden('ref_channelnorm_%s.npz' % tag)
    x, gout = torch.from_numpy(d['x']).to(DEV), torch.from_numpy(d['gout']).to(DEV)
    out = ops.channelnorm_fwd(x)
    assert np.array_equal(out.cpu().numpy(), d['out_fmad'])
    assert np.array_equal(ops.channelnorm_bwd(x, out, gout).cpu().numpy(), d['gin_fmad'])


@pytest.mark.gpu
@pytest.mark.parametrize('tag', CORR)
def test_correlation_kernels_equal_reference_kernels(tag):
    from imaginaire_b200 import ops
    d = H.load_golden('ref_correlation_%s.npz' % tag)
    cfg = [int(v) for v in d['cfg']]
    x1, x2, gout = (torch.from_numpy(d[k]).to(DEV) for k in ('x1', 'x2', 'gout'))
    out = ops.correlation_fwd(x1, x2, *cfg)
    assert np.array_equal(out.cpu().numpy(), d['out_fmad'])
    g1, g2 = ops.correlation_bwd(x1, x2, gout, *cfg)
    assert np.array_equal(g1.cpu().numpy(), d['g1_fmad']) and np.array_equal(g2.cpu().numpy(), d['g2_fmad'])
