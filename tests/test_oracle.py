"""The CPU oracle against the golden vectors produced by the reference itself
(oracle/make_golden.py).  CPU only; this is what pins the oracle."""
import numpy as np
import pytest
import torch

from oracle import spade_oracle as so
from oracle import warp_oracle as wo
from tests import helpers as H


@pytest.mark.parametrize('tag', ['a', 'b', 'c'])
def test_spade_norm_matches_reference(tag):
    d = H.load_golden('spade_norm_%s.npz' % tag)
    sd = H.split_sd(d, 'sd.')
    p = dict(kernel_size=int(d['kernel_size']), separate_projection=bool(d['separate_projection']),
             num_filters=int(d['num_filters']), norm_kind=str(d['norm_kind']))
    x, label = torch.from_numpy(d['x']), torch.from_numpy(d['label'])
    y = so.leaky_relu(so.spatially_adaptive_norm(sd, '', x, [label], p, bool(d['training'])))
    # same fp32 torch CPU arithmetic; only the op grouping differs -> a few ulp
    assert H.rel_err(y, d['y']) < 2e-6
    after = H.split_sd(d, 'sd_after.')
    for k in ('norm.running_mean', 'norm.running_var'):
        if k in after:
            assert H.rel_err(sd[k], after[k]) < 1e-6, k


@pytest.mark.parametrize('tag', ['a', 'b'])
def test_res2d_block_matches_reference(tag):
    d = H.load_golden('res2d_block_%s.npz' % tag)
    sd = H.split_sd(d, 'sd.')
    p = dict(kernel_size=5, separate_projection=True, num_filters=16, norm_kind=str(d['norm_kind']))
    x, label = torch.from_numpy(d['x']), torch.from_numpy(d['label'])
    y = so.spade_res_block(sd, '', x, label, p, bool(d['training']))
    assert H.rel_err(y, d['y']) < 1e-5
    after = H.split_sd(d, 'sd_after.')
    # spectral-norm power iteration and BN running statistics advance like the reference's
    for k, v in after.items():
        if k.endswith(('weight_u', 'weight_v', 'running_mean', 'running_var')):
            assert H.rel_err(sd[k], v) < 1e-5, k


def test_generator_unit_test_matches_reference():
    """BASELINE.json configs[0]: the unit-test SPADE generator, batch 1, on CPU."""
    d = H.load_golden('generator_unit_test.npz')
    cfg, net = H.build_generator('spade_unit_test.yaml')
    sd = net.state_dict()
    assert list(sd.keys()) == [str(k) for k in d['keys']]
    assert H.state_dict_digest(sd) == str(d['state_dict_sha256'])   # bit-identical seeded weights
    label, images = H.unit_test_generator_inputs()
    spec = so.generator_spec_from_cfg(cfg.gen, (256, 256))
    out = so.generator(sd, {'label': label, 'images': images}, spec, training=True,
                       eps_noise=torch.from_numpy(d['eps']))
    assert H.rel_err(out['mu'], d['mu']) < 1e-4
    assert H.rel_err(out['logvar'], d['logvar']) < 1e-4
    diff = (out['fake_images'] - torch.from_numpy(d['fake_images']).float()).abs()
    # golden is stored as fp16 (5e-4 resolution near 1); the fp32 paths agree to ~1e-4
    assert float(diff.max()) < 2e-3 and float(diff.mean()) < 3e-4


@pytest.mark.parametrize('tag', ['a', 'b'])
def test_gridsample_warp_oracle_bit_exact(tag):
    d = H.load_golden('warp_%s.npz' % tag)
    out = wo.gridsample_warp_forward(d['img'], d['flow'])
    assert np.array_equal(out, d['out'])


def test_gridsample_warp_oracle_large_case():
    d = H.load_golden('warp_c.npz')
    b, c, h, w = [int(v) for v in d['shape']]
    g = torch.Generator().manual_seed(30)
    img = torch.randn(b, c, h, w, generator=g)
    flow = H.synthetic_flow(31, b, h, w)
    out = wo.gridsample_warp_forward(img.numpy(), flow.numpy())
    assert np.array_equal(out[:, :, ::4, ::4], d['out_sub'])


def test_resample2d_oracle_known_answers():
    rng = np.random.RandomState(0)
    img = rng.randn(2, 3, 9, 11).astype(np.float32)
    zero = np.zeros((2, 2, 9, 11), np.float32)
    assert np.array_equal(wo.resample2d_forward(img, zero), img)          # zero flow = identity
    flow = zero.copy(); flow[:, 0] = 2; flow[:, 1] = -1                       # integer flow = exact shift
    ys = np.clip(np.arange(9) - 1, 0, 8); xs = np.clip(np.arange(11) + 2, 0, 10)
    assert np.array_equal(wo.resample2d_forward(img, flow), img[:, :, ys][:, :, :, xs])
    flow = zero.copy(); flow[:, 0] = 0.5                                      # half-pixel = average of neighbours
    out = wo.resample2d_forward(img, flow)
    exp = 0.5 * img + 0.5 * img[:, :, :, np.clip(np.arange(11) + 1, 0, 10)]
    assert np.allclose(out, exp, atol=1e-6)
    flow = zero.copy(); flow[:, 0] = 1000.0; flow[:, 1] = -1000.0            # far out of range = border clamp
    out = wo.resample2d_forward(img, flow)
    assert np.array_equal(out, np.broadcast_to(img[:, :, 0:1, 10:11], img.shape))


def test_resample2d_oracle_close_to_gridsample():
    """The two warps are the same function up to fp32 coordinate rounding."""
    d = H.load_golden('warp_a.npz')
    out = wo.resample2d_forward(d['img'], d['flow'])
    assert np.abs(out - d['out']).max() < 1e-3


def test_resample2d_oracle_backward_is_the_derivative():
    rng = np.random.RandomState(1)
    img = rng.randn(1, 2, 7, 8).astype(np.float32)
    flow = (rng.rand(1, 2, 7, 8).astype(np.float32) * 0.8 + 0.1)   # positive, fractional, interior
    flow[:, :, -2:, :] = 0.25; flow[:, :, :, -2:] = 0.25
    gout = rng.randn(1, 2, 7, 8).astype(np.float32)
    gimg, gflow = wo.resample2d_backward(img, flow, gout, trunc_alpha=True)
    # d/d image: <gout, d out/d img . e> by linearity
    e = rng.randn(*img.shape).astype(np.float32)
    lhs = float((gimg * e).sum())
    rhs = float((gout * wo.resample2d_forward(e, flow)).sum())
    assert abs(lhs - rhs) < 1e-3 * max(1.0, abs(rhs))
    # d/d flow by central differences at an interior pixel
    y, x = 2, 3
    for ch in (0, 1):
        fp, fm = flow.copy(), flow.copy()
        fp[0, ch, y, x] += 1e-2; fm[0, ch, y, x] -= 1e-2
        num = ((wo.resample2d_forward(img, fp) - wo.resample2d_forward(img, fm)) * gout).sum() / 2e-2
        assert abs(num - gflow[0, ch, y, x]) < 2e-2 * max(1.0, abs(num))


# ------------------------------------------------- FlowNet2 operators

def test_flownet_ops_oracle_known_answers():
    """The reference has neither tests nor a buildable extension for channelnorm / correlation
    ('parity unpinned'): the C restatement is held to the operators' mathematical definition,
    evaluated independently in float64 (forward) and by torch autograd (gradients)."""
    import numpy as np
    import torch
    from oracle import flownet_ops_oracle as fo
    rng = np.random.default_rng(0)
    x = rng.standard_normal((2, 5, 6, 7), dtype=np.float32)
    o = fo.channelnorm_forward(x)
    assert np.abs(o - np.sqrt((x.astype(np.float64) ** 2).sum(1, keepdims=True))).max() < 1e-6
    g = rng.standard_normal(o.shape, dtype=np.float32)
    assert np.abs(fo.channelnorm_backward(x, o, g) - g * x / (o + 1e-9)).max() < 1e-6

    x1 = rng.standard_normal((2, 40, 9, 11), dtype=np.float32)
    x2 = rng.standard_normal((2, 40, 9, 11), dtype=np.float32)
    pad, k, md, s1, s2 = 4, 1, 4, 1, 2
    assert fo.correlation_output_size(40, 9, 11, pad, k, md, s1, s2) == (25, 9, 11)
    assert fo.correlation_output_size(256, 48, 64, 20, 1, 20, 1, 2) == (441, 48, 64)     # FlowNetC
    out = fo.correlation_forward(x1, x2, pad, k, md, s1, s2)
    t1 = torch.tensor(x1, dtype=torch.float64, requires_grad=True)
    t2 = torch.tensor(x2, dtype=torch.float64, requires_grad=True)
    q1 = torch.nn.functional.pad(t1, (pad,) * 4)
    q2 = torch.nn.functional.pad(t2, (pad,) * 4)
    d, (hh, ww) = md // s2, out.shape[2:]
    ref = torch.stack([(q1[:, :, md:md + hh, md:md + ww] *
                        q2[:, :, md + tj * s2:md + tj * s2 + hh, md + ti * s2:md + ti * s2 + ww]).sum(1) / 40
                       for tj in range(-d, d + 1) for ti in range(-d, d + 1)], 1)
    assert np.abs(out - ref.detach().numpy()).max() < 1e-6
    go = rng.standard_normal(out.shape, dtype=np.float32)
    ref.backward(torch.tensor(go, dtype=torch.float64))
    g1, g2 = fo.correlation_backward(x1, x2, go, pad, k, md, s1, s2)
    assert np.abs(g1 - t1.grad.numpy()).max() < 1e-6
    assert np.abs(g2 - t2.grad.numpy()).max() < 1e-6
