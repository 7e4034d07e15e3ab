"""GPU parity of the C-ABI kernels against the CPU oracle (oracle/) and the golden vectors.
Bit-exact where the arithmetic is reproducible (warp forwards, resizes, packs), stated
tolerances elsewhere."""
import os

import numpy as np
import pytest
import torch
import torch.nn.functional as F

from oracle import spade_oracle as so
from oracle import warp_oracle as wo
from tests import helpers as H

pytestmark = pytest.mark.gpu
DEV = 'cuda:0'


def nhwc(t):
    return t.to(DEV).bfloat16().contiguous(memory_format=torch.channels_last)


# ------------------------------------------------------------------ warps

@pytest.mark.parametrize('tag', ['a', 'b'])
def test_gridsample_warp_bit_exact_vs_reference_golden(tag):
    from imaginaire_b200.model_utils.fs_vid2vid import resample
    d = H.load_golden('warp_%s.npz' % tag)
    out = resample(torch.from_numpy(d['img']).to(DEV), torch.from_numpy(d['flow']).to(DEV))
    assert np.array_equal(out.cpu().numpy(), d['out'])


def test_gridsample_warp_full_size_512x1024():
    """BASELINE.json configs[3] shape: 512 x 1024 frames; 1e-5 relative bound (measured: bit-exact)."""
    from imaginaire_b200 import ops
    g = torch.Generator().manual_seed(7)
    img = torch.randn(2, 3, 512, 1024, generator=g)
    flow = H.synthetic_flow(8, 2, 512, 1024)
    out = ops.gridsample_warp_fwd(img.to(DEV), flow.to(DEV)).cpu().numpy()
    ref = wo.gridsample_warp_forward(img.numpy(), flow.numpy())
    assert np.abs(out - ref).max() <= 1e-5 * np.abs(ref).max()
    assert np.array_equal(out, ref)


@pytest.mark.parametrize('shape', [(2, 3, 64, 128), (1, 5, 33, 47), (2, 3, 512, 1024), (1, 1, 1, 1), (3, 2, 1, 9)])
def test_resample2d_forward_bit_exact_vs_oracle(shape):
    from imaginaire_b200.third_party.resample2d import Resample2d
    b, c, h, w = shape
    g = torch.Generator().manual_seed(11)
    img = torch.randn(b, c, h, w, generator=g)
    flow = H.synthetic_flow(12, b, h, w) if h > 4 else torch.randn(b, 2, h, w, generator=g) * 3
    out = Resample2d()(img.to(DEV), flow.to(DEV)).cpu().numpy()
    ref = wo.resample2d_forward(img.numpy(), flow.numpy())
    assert np.abs(out - ref).max() <= 1e-5 * max(np.abs(ref).max(), 1e-30)   # the stated tolerance
    assert np.array_equal(out, ref)                                          # and in fact bit-exact


@pytest.mark.parametrize('shape,scale', [((1, 16, 512, 1024), 8.0), ((2, 9, 500, 1028), 8.0), ((1, 16, 512, 1024), 24.0),
                                         ((1, 16, 512, 1024), 0.0)])
@pytest.mark.parametrize('which', ['resample2d', 'gridsample'])
def test_warp_forward_through_the_shared_memory_window_is_bit_exact(shape, scale, which, monkeypatch):
    """Large launches read their taps from a TMA-staged source window (csrc/warp_window.cu): same arithmetic, so
    still bit-exact.  The flows cover all three box sizes (scale 0 / 8 / 24 px), ragged tiles (H = 500), and tiles
    whose taps fit no box (the out-of-range rows of the SURVEY flow set) that fall back to per-pixel gathers."""
    from imaginaire_b200 import ops
    from imaginaire_b200.model_utils.fs_vid2vid import resample
    monkeypatch.setenv('IMX_WARPWIN_R2D', '1')     # resample2d keeps the per-pixel kernel by default; test both
    b, c, h, w = shape
    g = torch.Generator().manual_seed(21)
    img = torch.randn(b, c, h, w, generator=g)
    flow = H.synthetic_flow(22, b, h, w, scale=scale)
    if which == 'resample2d':
        out = ops.resample2d_fwd(img.to(DEV), flow.to(DEV)).cpu().numpy()
        ref = wo.resample2d_forward(img.numpy(), flow.numpy())
    else:
        out = resample(img.to(DEV), flow.to(DEV)).cpu().numpy()
        ref = wo.gridsample_warp_forward(img.numpy(), flow.numpy())
    assert np.array_equal(out, ref)


def test_resample2d_empty_and_invalid():
    from imaginaire_b200 import ops, _C
    out = ops.resample2d_fwd(torch.zeros(0, 3, 8, 8, device=DEV), torch.zeros(0, 2, 8, 8, device=DEV))
    assert out.shape == (0, 3, 8, 8)
    with pytest.raises(_C.ImxError, match='kernel_size'):
        ops.resample2d_fwd(torch.zeros(1, 3, 8, 8, device=DEV), torch.zeros(1, 2, 8, 8, device=DEV), kernel_size=3)
    with pytest.raises(RuntimeError):
        ops.resample2d_fwd(torch.zeros(1, 3, 8, 8, device=DEV).half(), torch.zeros(1, 2, 8, 8, device=DEV))


@pytest.mark.parametrize('trunc', [True, False])
def test_resample2d_backward_vs_oracle(trunc):
    from imaginaire_b200 import ops
    g = torch.Generator().manual_seed(13)
    img = torch.randn(2, 4, 40, 56, generator=g)
    flow = H.synthetic_flow(14, 2, 40, 56, scale=4.0)
    gout = torch.randn(2, 4, 40, 56, generator=g)
    gi, gf = ops.resample2d_bwd(img.to(DEV), flow.to(DEV), gout.to(DEV), trunc_alpha=trunc)
    ri, rf = wo.resample2d_backward(img.numpy(), flow.numpy(), gout.numpy(), trunc_alpha=trunc)
    assert np.array_equal(gf.cpu().numpy(), rf)                 # per-pixel, same op order: bit-exact
    assert H.rel_err(gi.cpu(), ri) < 1e-5                        # atomics: summation order differs


def test_resample2d_autograd_function():
    from imaginaire_b200.third_party.resample2d import Resample2dFunction
    g = torch.Generator().manual_seed(15)
    img = torch.randn(1, 3, 16, 20, generator=g).to(DEV).requires_grad_(True)
    flow = (torch.rand(1, 2, 16, 20, generator=g) * 2).to(DEV).requires_grad_(True)
    out = Resample2dFunction.apply(img, flow, 1)
    gout = torch.randn(1, 3, 16, 20, generator=g)
    out.backward(gout.to(DEV))
    ri, rf = wo.resample2d_backward(img.detach().cpu().numpy(), flow.detach().cpu().numpy(), gout.numpy(), True)
    assert H.rel_err(img.grad.cpu(), ri) < 1e-5 and np.array_equal(flow.grad.cpu().numpy(), rf)


def test_gridsample_warp_backward_vs_torch_autograd():
    from imaginaire_b200.model_utils.fs_vid2vid import resample
    g = torch.Generator().manual_seed(16)
    img = torch.randn(2, 3, 48, 64, generator=g)
    flow = H.synthetic_flow(17, 2, 48, 64, scale=5.0)
    gout = torch.randn(2, 3, 48, 64, generator=g)
    a = img.to(DEV).requires_grad_(True); f = flow.to(DEV).requires_grad_(True)
    resample(a, f).backward(gout.to(DEV))
    # the reference's resample() written out with torch CPU ops (fs_vid2vid.py:14-39)
    ic = img.clone().requires_grad_(True); fc = flow.clone().requires_grad_(True)
    h, w = 48, 64
    gx = torch.linspace(-1, 1, w).view(1, 1, 1, w).expand(2, 1, h, w)
    gy = torch.linspace(-1, 1, h).view(1, 1, h, 1).expand(2, 1, h, w)
    fl = torch.cat([fc[:, 0:1] / ((w - 1.0) / 2.0), fc[:, 1:2] / ((h - 1.0) / 2.0)], 1)
    F.grid_sample(ic, (torch.cat([gx, gy], 1) + fl).permute(0, 2, 3, 1), mode='bilinear', padding_mode='border',
                  align_corners=True).backward(gout)
    assert H.rel_err(a.grad.cpu(), ic.grad) < 1e-5
    assert H.rel_err(f.grad.cpu(), fc.grad) < 1e-4


# ------------------------------------------------------- layout / resize

@pytest.mark.parametrize('size', [(64, 96), (16, 24), (5, 7), (1, 1)])
def test_nearest_resize_exact_vs_oracle(size):
    from imaginaire_b200 import ops
    label = H.onehot_label(21, 2, 35, 64, 96, extra_uniform=1)
    out = ops.nearest_resize_to_nhwc(label.to(DEV), *size)
    ref = so.nearest_resize(label, size).bfloat16().float()
    assert out.shape[1] == 64
    assert torch.equal(out[:, :36].float().cpu(), ref)
    assert float(out[:, 36:].float().abs().max()) == 0.0


def test_upsample2x_exact_and_backward():
    from imaginaire_b200 import functional as NF
    x = nhwc(torch.randn(2, 64, 6, 10)).requires_grad_(True)
    y = NF.upsample2x(x)
    assert torch.equal(y.float().cpu(), so.upsample2x(x.detach().float().cpu()))
    g = nhwc(torch.randn(2, 64, 12, 20))
    y.backward(g)
    ref = F.avg_pool2d(g.float(), 2) * 4
    assert H.rel_err(x.grad.float(), ref) < 8e-3   # bf16 rounding of the 4-sum


# ------------------------------------------------------------ convolution

CONV_CASES = [
    # N, Cin, Cout, H, W, k, stride, pad
    (1, 64, 32, 8, 16, 1, 1, 0),
    (2, 64, 64, 20, 24, 3, 1, 1),          # ragged tiles
    (1, 40, 72, 13, 11, 5, 1, 2),          # channel counts that are not multiples of 64
    (2, 192, 64, 16, 16, 5, 1, 2),         # label -> hidden
    (2, 256, 512, 16, 32, 3, 1, 1),
    (1, 256, 3, 64, 64, 5, 1, 2),          # to-RGB
    (4, 512, 1024, 32, 64, 3, 1, 1),       # > 148 tiles: persistent loop, both TMEM stages
    (2, 64, 64, 32, 32, 3, 2, 1),          # stride 2 (style encoder)
    (2, 64, 128, 33, 47, 4, 2, 1),         # discriminator-style k4 s2, odd size
    (1, 8, 16, 7, 5, 3, 1, 1),             # smallest
]


@pytest.mark.parametrize('case', CONV_CASES)
def test_conv2d_fprop_vs_oracle(case):
    """fp32 CPU convolution of the bf16-rounded operands is the oracle; the kernel accumulates
    in fp32 and rounds the output to bf16 once: |err| <= 2^-8 relative to the output scale."""
    from imaginaire_b200 import ops
    n, cin, cout, h, w, k, s, p = case
    g = torch.Generator().manual_seed(31)
    x = torch.randn(n, cin, h, w, generator=g).bfloat16().float()
    wt = (torch.randn(cout, cin, k, k, generator=g) / (cin * k * k) ** 0.5).bfloat16().float()
    b = torch.randn(cout, generator=g)
    cin_p = ops.round_up(cin, 8)
    xp = torch.zeros(n, cin_p, h, w); xp[:, :cin] = x
    y = ops.conv2d_fprop(nhwc(xp), ops.pack_conv_weight(wt.to(DEV), cin_p=cin_p), cout, k, k, s, p, bias=b.to(DEV))
    ref = F.conv2d(x, wt, b, stride=s, padding=p)
    assert y.shape[1] == ops.round_up(cout, 8)
    assert H.rel_err(y[:, :cout].float().cpu(), ref) < 4e-3
    assert float(y[:, cout:].float().abs().max() if y.shape[1] > cout else 0.0) == 0.0


# The shapes the timed C3 iteration runs (batch reduced where the tile schedule does not depend on it): 256-pixel
# tiles (MT = 2) with more than one wave, BN = 256 over several waves at 128 x 256, 2048-channel K loops, the 5x5
# projections on 35 -> 128 -> C channels, the discriminator's first k4 s2 layer, the to-RGB layer.
BENCH_CONV_CASES = [
    # N, Cin, Cout, H, W, k, stride, pad
    (2, 512, 512, 128, 256, 3, 1, 1),      # BN = 256, 1024 tiles: many waves of the persistent loop
    (2, 512, 256, 128, 256, 3, 1, 1),
    (2, 128, 512, 128, 256, 5, 1, 2),      # gamma / beta projection at the largest resolution
    (8, 40, 128, 64, 128, 5, 1, 2),        # label -> hidden (35 labels padded to 40), BN = 128 with MT = 2
    (8, 128, 1024, 32, 64, 5, 1, 2),
    (8, 2048, 2048, 16, 32, 3, 1, 1),      # 2048-channel K loop, split-K territory (128 tiles)
    (8, 2048, 1024, 32, 64, 1, 1, 0),      # learned shortcut
    (4, 40, 128, 256, 512, 4, 2, 1),       # discriminator layer 0
    (2, 256, 3, 256, 512, 5, 1, 2),        # to-RGB (shift-add path is taken by the module; this is the kernel)
]


@pytest.mark.parametrize('case', BENCH_CONV_CASES)
def test_conv2d_fprop_at_the_timed_shapes(case):
    """fprop against the fp32 CPU convolution on bf16-rounded operands, every output element, at the layer shapes of
    the timed iteration (tests/test_gpu_kernels.py::CONV_CASES stops at 4 x 512 -> 1024 @ 32 x 64)."""
    from imaginaire_b200 import ops
    n, cin, cout, h, w, k, s, p = case
    g = torch.Generator().manual_seed(131)
    x = torch.randn(n, cin, h, w, generator=g).bfloat16().float()
    wt = (torch.randn(cout, cin, k, k, generator=g) / (cin * k * k) ** 0.5).bfloat16().float()
    b = torch.randn(cout, generator=g)
    y = ops.conv2d_fprop(nhwc(x), ops.pack_conv_weight(wt.to(DEV), cin_p=cin), cout, k, k, s, p, bias=b.to(DEV))
    torch.set_num_threads(max(1, (os.cpu_count() or 1)))
    ref = F.conv2d(x, wt, b, stride=s, padding=p)
    assert H.rel_err(y[:, :cout].float().cpu(), ref) < 4e-3


@pytest.mark.parametrize('case', [(8, 2048, 2048, 16, 32, 3, 1, 1),     # 587 TF/s shape: few pixels, many taps x tiles
                                  (8, 128, 2048, 16, 32, 5, 1, 2),
                                  (4, 64, 128, 256, 512, 5, 1, 2),      # many pixels: deep split over the pixel axis
                                  (2, 512, 512, 128, 256, 3, 1, 1),
                                  (8, 1024, 512, 64, 128, 1, 1, 0)])
def test_conv2d_wgrad_at_the_timed_shapes(case):
    """The split planner at C3 sizes: filter gradient against torch's CPU computation (exact bf16 products, fp32
    accumulation on both sides; the GPU sums its split-K partials in another order)."""
    from imaginaire_b200 import ops
    n, cin, cout, h, w, k, s, p = case
    g = torch.Generator().manual_seed(133)
    ho, wo_ = (h + 2 * p - k) // s + 1, (w + 2 * p - k) // s + 1
    x = torch.randn(n, cin, h, w, generator=g).bfloat16().float()
    dy = (torch.randn(n, cout, ho, wo_, generator=g) / (n * ho * wo_) ** 0.5).bfloat16().float()
    dw = ops.conv2d_wgrad(nhwc(x), nhwc(dy), cin, cout, k, k, s, p)
    torch.set_num_threads(max(1, (os.cpu_count() or 1)))
    ref = torch.nn.grad.conv2d_weight(x, (cout, cin, k, k), dy, stride=s, padding=p)
    assert H.rel_err(dw.cpu(), ref) < 2e-4


def test_conv2d_epilogues_vs_oracle():
    from imaginaire_b200 import ops
    g = torch.Generator().manual_seed(32)
    n, cin, cout, h, w = 2, 128, 128, 16, 32
    x = torch.randn(n, cin, h, w, generator=g).bfloat16().float()
    wt = (torch.randn(cout, cin, 3, 3, generator=g) / (cin * 9) ** 0.5).bfloat16().float()
    b = torch.randn(cout, generator=g)
    res = torch.randn(n, cout, h, w, generator=g).bfloat16().float()
    wpk = ops.pack_conv_weight(wt.to(DEV))
    base = F.conv2d(x, wt, b, padding=1)
    y = ops.conv2d_fprop(nhwc(x), wpk, cout, 3, 3, 1, 1, bias=b.to(DEV), flags=ops.EPI_LRELU)
    assert H.rel_err(y.float().cpu(), so.leaky_relu(base)) < 4e-3
    y = ops.conv2d_fprop(nhwc(x), wpk, cout, 3, 3, 1, 1, bias=b.to(DEV), flags=ops.EPI_RELU)
    assert H.rel_err(y.float().cpu(), F.relu(base)) < 4e-3
    y = ops.conv2d_fprop(nhwc(x), wpk, cout, 3, 3, 1, 1, bias=b.to(DEV), residual=nhwc(res))
    assert H.rel_err(y.float().cpu(), base + res) < 4e-3
    y = ops.conv2d_fprop(nhwc(x), wpk, cout, 3, 3, 1, 1, bias=b.to(DEV), flags=ops.EPI_TANH)
    assert H.rel_err(y.float().cpu(), torch.tanh(base)) < 4e-3
    y = ops.conv2d_fprop(nhwc(x), wpk, cout, 3, 3, 1, 1, bias=b.to(DEV), out_f32=True)
    assert H.rel_err(y.float().cpu(), base) < 1e-5
    for inst in (False, True):
        st = torch.zeros(2, n if inst else 1, cout, device=DEV)
        y = ops.conv2d_fprop(nhwc(x), wpk, cout, 3, 3, 1, 1, bias=b.to(DEV), stats=st)
        yq = y.float().cpu()
        dims = (2, 3) if inst else (0, 2, 3)
        assert H.rel_err(st[0].cpu().flatten(), yq.sum(dims).flatten()) < 1e-4
        assert H.rel_err(st[1].cpu().flatten(), (yq * yq).sum(dims).flatten()) < 1e-4


@pytest.mark.parametrize('case', [(8, 2048, 128, 16, 32, 5, 1, 2), (8, 512, 512, 16, 32, 4, 2, 1), (2, 256, 40, 9, 13, 3, 1, 1)])
def test_conv2d_fprop_split_k_matches_plain_tiling(case):
    """Layers with fewer pixel tiles than SMs split the K loop over CTAs (fp32 partial sums in a
    workspace + finishing kernel).  Same operands through the plain entry point: the only
    difference is the fp32 summation order, far below the bf16 rounding of the output."""
    from imaginaire_b200 import ops, _C
    n, cin, cout, h, w, k, s, p = case
    g = torch.Generator().manual_seed(35)
    x = nhwc(torch.randn(n, cin, h, w, generator=g))
    wpk = ops.pack_conv_weight((torch.randn(cout, cin, k, k, generator=g) / (cin * k * k) ** 0.5).to(DEV))
    b = torch.randn(cout, generator=g).to(DEV)
    res_shape = (n, ops.round_up(cout, 8), (h + 2 * p - k) // s + 1, (w + 2 * p - k) // s + 1)
    res = nhwc(torch.randn(*res_shape, generator=g))
    lib = _C.lib()
    flags = ops.EPI_BIAS | ops.EPI_LRELU | ops.EPI_RESIDUAL
    need = lib.raw('imx_conv2d_fprop_workspace_bytes')(n, h, w, cin, cout, res_shape[1], k, k, s, p, flags)
    assert need == n * res_shape[2] * res_shape[3] * res_shape[1] * 4      # split-K chosen for these shapes
    y_split = ops.conv2d_fprop(x, wpk, cout, k, k, s, p, bias=b, residual=res, flags=ops.EPI_LRELU)
    y_plain = ops.empty_nhwc(*res_shape, x.device)
    lib.call('imx_conv2d_igemm_fprop_bf16', x.data_ptr(), wpk.data_ptr(), b.data_ptr(), res.data_ptr(), y_plain.data_ptr(),
             None, n, h, w, cin, cout, res_shape[1], k, k, s, p, flags, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    # two bf16-rounded results of differently ordered fp32 sums differ by at most one bf16 ulp (2^-7 relative)
    assert H.rel_err(y_split.float().cpu(), y_plain.float().cpu()) < 1e-2
    assert float((y_split.float() - y_plain.float()).abs().mean() / y_plain.float().abs().mean()) < 2e-3
    # a too-small workspace is refused, not overrun
    with pytest.raises(_C.ImxError):
        ws = torch.empty(16, dtype=torch.uint8, device=DEV)
        lib.call('imx_conv2d_igemm_fprop_splitk_bf16', x.data_ptr(), wpk.data_ptr(), b.data_ptr(), res.data_ptr(),
                 y_plain.data_ptr(), None, ws.data_ptr(), 16, n, h, w, cin, cout, res_shape[1], k, k, s, p, flags,
                 torch.cuda.current_stream().cuda_stream)


@pytest.mark.parametrize('case', [(1, 64, 128, 8, 8, 1, 1, 0), (2, 64, 64, 16, 32, 3, 1, 1), (1, 40, 72, 13, 11, 5, 1, 2),
                                  (2, 64, 512, 32, 64, 5, 1, 2), (2, 64, 128, 32, 32, 4, 2, 1),
                                  (4, 256, 256, 32, 64, 3, 1, 1)])
def test_conv2d_wgrad_vs_oracle(case):
    from imaginaire_b200 import ops
    n, cin, cout, h, w, k, s, p = case
    g = torch.Generator().manual_seed(33)
    ho, wo_ = (h + 2 * p - k) // s + 1, (w + 2 * p - k) // s + 1
    x = torch.randn(n, cin, h, w, generator=g).bfloat16().float()
    dy = torch.randn(n, cout, ho, wo_, generator=g).bfloat16().float()
    cin_p, cout_p = ops.round_up(cin, 8), ops.round_up(cout, 8)
    xp = torch.zeros(n, cin_p, h, w); xp[:, :cin] = x
    dp = torch.zeros(n, cout_p, ho, wo_); dp[:, :cout] = dy
    dw = ops.conv2d_wgrad(nhwc(xp), nhwc(dp), cin, cout, k, k, s, p)
    ref = torch.nn.grad.conv2d_weight(x, (cout, cin, k, k), dy, stride=s, padding=p)
    assert H.rel_err(dw.cpu(), ref) < 1e-4      # exact bf16 products, fp32 accumulation both sides


@pytest.mark.parametrize('case', [(2, 64, 64, 16, 32, 3, 1, 1), (1, 40, 72, 13, 11, 5, 1, 2), (2, 64, 128, 32, 32, 3, 2, 1),
                                  (2, 32, 64, 33, 47, 4, 2, 1), (1, 40, 128, 64, 128, 4, 2, 1), (2, 128, 256, 30, 62, 4, 2, 2),
                                  (3, 72, 40, 17, 9, 3, 2, 1), (1, 512, 512, 16, 32, 4, 2, 1), (2, 64, 64, 21, 35, 5, 2, 2),
                                  (2, 64, 96, 16, 16, 2, 2, 0)])
def test_conv2d_dgrad_vs_oracle(case):
    from imaginaire_b200 import ops, functional as NF
    n, cin, cout, h, w, k, s, p = case
    g = torch.Generator().manual_seed(34)
    ho, wo_ = (h + 2 * p - k) // s + 1, (w + 2 * p - k) // s + 1
    wt = (torch.randn(cout, cin, k, k, generator=g) / (cout * k * k) ** 0.5).bfloat16().float()
    dy = torch.randn(n, cout, ho, wo_, generator=g).bfloat16().float()
    cin_p, cout_p = ops.round_up(cin, 8), ops.round_up(cout, 8)
    dp = torch.zeros(n, cout_p, ho, wo_); dp[:, :cout] = dy
    dx = NF.conv2d_dgrad(nhwc(dp), wt.to(DEV), s, p, h, w, cin_p)
    ref = torch.nn.grad.conv2d_input((n, cin, h, w), wt, dy, stride=s, padding=p)
    assert dx.shape == (n, cin_p, h, w)
    assert H.rel_err(dx[:, :cin].float().cpu(), ref) < 4e-3
    if s == 2:
        # the parity-class form (four stride-1 convolutions) and the zero-insertion form are the same sums
        os.environ['IMX_DGRAD_S2'] = '0'
        try:
            dz = NF.conv2d_dgrad(nhwc(dp), wt.to(DEV), s, p, h, w, cin_p)
        finally:
            del os.environ['IMX_DGRAD_S2']
        assert H.rel_err(dx[:, :cin].float().cpu(), dz[:, :cin].float().cpu()) < 4e-3
        # a Parameter with 1 / sigma: sub-filters packed unscaled (cached), the scale applied in the epilogue
        wp = torch.nn.Parameter(wt.to(DEV))
        inv_sigma = torch.tensor([0.61], device=DEV)
        for _ in range(2):          # second call: cache hit
            ds = NF.conv2d_dgrad(nhwc(dp), wp, s, p, h, w, cin_p, inv_sigma=inv_sigma)
        assert H.rel_err(ds[:, :cin].float().cpu(), ref * 0.61) < 4e-3


@pytest.mark.parametrize('case', [(2, 3, 16, 20, 33, 3, 1, 1),      # few input channels (image -> first layer)
                                  (2, 3, 32, 32, 48, 3, 2, 1),      # ... stride 2; dgrad = zero insertion + tap sum
                                  (1, 3, 24, 17, 19, 5, 1, 2),
                                  (2, 40, 3, 21, 37, 5, 1, 2),      # few output channels (to-RGB)
                                  (2, 256, 3, 32, 64, 5, 1, 2),
                                  (1, 64, 1, 16, 16, 3, 1, 1)])
def test_small_channel_convs_forward_and_gradients(case):
    """Layers with <= 8 channels on one side run as im2col / per-tap-product GEMMs plus a
    memory-bound pass (csrc/small_conv.cu).  Forward, dx, dW and dbias against torch autograd
    on the same bf16-rounded operands."""
    from imaginaire_b200 import ops, functional as NF
    n, cin, cout, h, w, k, s, p = case
    g = torch.Generator().manual_seed(36)
    x = torch.randn(n, cin, h, w, generator=g).bfloat16().float()
    wt = (torch.randn(cout, cin, k, k, generator=g) / (cin * k * k) ** 0.5).bfloat16().float()
    b = torch.randn(cout, generator=g)
    cin_p = ops.round_up(cin, 8)
    xp = torch.zeros(n, cin_p, h, w); xp[:, :cin] = x
    xd = nhwc(xp).requires_grad_(True)
    wd = wt.to(DEV).requires_grad_(True)
    bd = b.to(DEV).requires_grad_(True)
    y = NF.conv2d(xd, wd, bd, stride=s, pad=p)
    xr, wr, br = x.clone().requires_grad_(True), wt.clone().requires_grad_(True), b.clone().requires_grad_(True)
    ref = F.conv2d(xr, wr, br, stride=s, padding=p)
    assert y.shape == (n, ops.round_up(cout, 8)) + tuple(ref.shape[2:])
    assert H.rel_err(y[:, :cout].float().cpu(), ref) < 6e-3
    assert float(y[:, cout:].float().abs().max() if y.shape[1] > cout else 0.0) == 0.0
    dy = torch.randn(ref.shape, generator=g).bfloat16().float()
    dyp = torch.zeros(y.shape); dyp[:, :cout] = dy
    y.backward(nhwc(dyp))
    ref.backward(dy)
    assert H.rel_err(xd.grad[:, :cin].float().cpu(), xr.grad) < 6e-3
    assert H.rel_err(wd.grad.cpu(), wr.grad) < 6e-3
    assert H.rel_err(bd.grad.cpu(), br.grad) < 1e-4


# ------------------------------------------------- statistics / modulation

@pytest.mark.parametrize('shape', [(2, 64, 16, 32), (3, 512, 9, 7), (1, 8, 1, 1), (2, 2048, 4, 8)])
@pytest.mark.parametrize('act', [1, 2])
def test_act_backward_with_bias_grad_exact(shape, act):
    """dx = dy * act'(y) and its per-channel sum in one pass; exact against the same arithmetic in torch."""
    from imaginaire_b200 import ops
    g = torch.Generator().manual_seed(41)
    y = nhwc(torch.randn(*shape, generator=g))
    dy = nhwc(torch.randn(*shape, generator=g))
    dx, bg = ops.act_bwd_bias(dy, y, act, True)
    slope = 0.2 if act == 1 else 0.0
    ref = torch.where(y.float() > 0, dy.float(), dy.float() * slope).bfloat16()
    assert torch.equal(dx.cpu(), ref.cpu())
    assert H.rel_err(bg.cpu(), ref.float().sum((0, 2, 3)).cpu()) < 1e-5
    dx2, none = ops.act_bwd_bias(dy, y, act, False)
    assert none is None and torch.equal(dx2, dx)


@pytest.mark.parametrize('shape', [(2, 64, 16, 32), (3, 512, 8, 8), (2, 2048, 4, 8), (1, 40, 5, 7), (2, 4096, 2, 2)])
@pytest.mark.parametrize('inst', [False, True])
def test_bn_stats_and_finalize_vs_oracle(shape, inst):
    from imaginaire_b200 import ops
    n, c, h, w = shape
    g = torch.Generator().manual_seed(41)
    x = (torch.randn(n, c, h, w, generator=g) * 1.7 + 0.3).bfloat16().float()
    sums = ops.bn_stats(nhwc(x), inst)
    dims = (2, 3) if inst else (0, 2, 3)
    assert H.rel_err(sums[0].cpu().flatten(), x.sum(dims).flatten()) < 1e-5
    assert H.rel_err(sums[1].cpu().flatten(), (x * x).sum(dims).flatten()) < 1e-5
    rm, rv = torch.zeros(c, device=DEV), torch.ones(c, device=DEV)
    cnt = h * w if inst else n * h * w
    mean, rstd = ops.bn_finalize(sums, cnt, 1e-5, None if inst else rm, None if inst else rv, 0.1)
    assert H.rel_err(mean.cpu(), x.mean(dims).flatten()) < 1e-5
    assert H.rel_err(rstd.cpu(), (x.var(dims, unbiased=False) + 1e-5).rsqrt().flatten()) < 1e-4
    if not inst and cnt > 1:
        # oracle: running = 0.9 * running + 0.1 * {mean, unbiased var}  (spade_oracle.inner_norm)
        assert H.rel_err(rv.cpu(), 0.9 + 0.1 * x.var(dims, unbiased=True)) < 1e-4
        assert H.rel_err(rm.cpu(), 0.1 * x.mean(dims)) < 1e-4


@pytest.mark.parametrize('inst', [False, True])
def test_bn_stats_with_a_large_mean_and_many_elements(inst):
    """variance by E[x^2] - E[x]^2 in fp32 cancels when |mean| >> sigma and the count is large.
    1 M elements per channel (262 k per instance), |mean| / sigma up to 30: mean and rstd against float64."""
    from imaginaire_b200 import ops
    g = torch.Generator().manual_seed(52)
    n, c, h, w = 4, 16, 512, 512
    ratio = torch.tensor([0.0, 1.0, 3.0, 10.0, 30.0, -30.0, 20.0, 5.0] * 2)
    x = (torch.randn(n, c, h, w, generator=g) + ratio.view(1, c, 1, 1)).bfloat16().float()
    sums = ops.bn_stats(nhwc(x), inst)
    count = h * w if inst else n * h * w
    mean, rstd = ops.bn_finalize(sums, count, 1e-5)
    xd = x.double()
    dims = (2, 3) if inst else (0, 2, 3)
    m_ref = xd.mean(dims).reshape(-1)
    r_ref = torch.rsqrt(xd.var(dims, unbiased=False) + 1e-5).reshape(-1)
    assert H.rel_err(mean.cpu(), m_ref) < 1e-5
    err = float(((rstd.cpu().double() - r_ref) / r_ref).abs().max())
    print('worst relative rstd error %.2e at |mean|/sigma up to 30, %d elements per statistic' % (err, count))
    assert err < 2e-3          # a bf16 activation carries 4e-3; the cancellation must stay below that


def test_spade_gb_fused_equals_materialised():
    """The projection-conv-applies-the-modulation path and the materialised gamma/beta path
    are the same function (bf16 rounding of gamma/beta aside)."""
    from imaginaire_b200 import ops
    g = torch.Generator().manual_seed(42)
    n, ch, c, h, w, k = 2, 128, 512, 16, 32, 5
    hid = F.relu(torch.randn(n, ch, h, w, generator=g)).bfloat16().float()
    wg = (torch.randn(c, ch, k, k, generator=g) / (ch * k * k) ** 0.5)
    wb = (torch.randn(c, ch, k, k, generator=g) / (ch * k * k) ** 0.5)
    bg, bb = torch.randn(c, generator=g) * 0.1, torch.randn(c, generator=g) * 0.1
    x = (torch.randn(n, c, h, w, generator=g) * 1.5 + 0.2).bfloat16().float()
    nw, nb = torch.rand(c, generator=g) + 0.5, torch.randn(c, generator=g) * 0.3
    xn = nhwc(x)
    sums = ops.bn_stats(xn, False)
    mean, rstd = ops.bn_finalize(sums, n * h * w, 1e-5)
    y = ops.spade_gb_conv_modulate(nhwc(hid), ops.pack_spade_gb_weight(wg.to(DEV), wb.to(DEV)), bg.to(DEV), bb.to(DEV),
                                   xn, mean, rstd, nw.to(DEV), nb.to(DEV), c, k, k, k // 2, False, ops.ACT_LRELU)
    # oracle: the same formula on CPU in fp32
    gamma = F.conv2d(hid, wg.bfloat16().float(), bg, padding=k // 2)
    beta = F.conv2d(hid, wb.bfloat16().float(), bb, padding=k // 2)
    m = x.mean((0, 2, 3), keepdim=True); v = x.var((0, 2, 3), unbiased=False, keepdim=True)
    xh = (x - m) / torch.sqrt(v + 1e-5) * nw.view(1, -1, 1, 1) + nb.view(1, -1, 1, 1)
    ref = so.leaky_relu(xh * (1 + gamma) + beta)
    assert H.rel_err(y.float().cpu(), ref) < 6e-3


# ------------------------------------------------------------ spectral norm

@pytest.mark.parametrize('shape', [(64, 35, 5, 5), (128, 128, 3, 3), (2048, 256, 3, 3), (8, 130, 1, 1), (3, 256, 5, 5)])
@pytest.mark.parametrize('training', [True, False])
def test_spectral_norm_sigma_vs_oracle(shape, training):
    """Power iteration + sigma against the oracle's restatement of torch's spectral_norm."""
    from imaginaire_b200 import ops
    g = torch.Generator().manual_seed(51)
    w = torch.randn(*shape, generator=g) * 0.1
    u = F.normalize(torch.randn(shape[0], generator=g), dim=0)
    v = F.normalize(torch.randn(w.numel() // shape[0], generator=g), dim=0)
    sd = {'weight_orig': w.clone(), 'weight_u': u.clone(), 'weight_v': v.clone()}
    w_eff = so.effective_weight(sd, '', training)            # updates sd's u, v when training
    sigma_ref = float((w.flatten()[0] / w_eff.flatten()[0]))
    ug, vg = u.to(DEV), v.to(DEV)
    inv_sigma, uc, vc = ops.spectral_norm_sigma(w.to(DEV), ug, vg, 1 if training else 0, 1e-12)
    assert abs(1.0 / float(inv_sigma) - sigma_ref) < 1e-4 * abs(sigma_ref)
    assert H.rel_err(ug.cpu(), sd['weight_u']) < 1e-4 and H.rel_err(vg.cpu(), sd['weight_v']) < 1e-4
    assert torch.equal(uc, ug) and torch.equal(vc, vg)        # the copies autograd keeps


def test_spectral_norm_weight_grad_vs_autograd():
    from imaginaire_b200 import ops
    g = torch.Generator().manual_seed(52)
    w = (torch.randn(96, 40, 3, 3, generator=g) * 0.1).requires_grad_(True)
    u = F.normalize(torch.randn(96, generator=g), dim=0)
    v = F.normalize(torch.randn(360, generator=g), dim=0)
    grad_eff = torch.randn(96, 40, 3, 3, generator=g)
    sigma = torch.dot(u, torch.mv(w.reshape(96, -1), v))
    (w / sigma).backward(grad_eff)
    out = ops.spectral_norm_weight_grad_(grad_eff.clone().to(DEV), w.detach().to(DEV), u.to(DEV), v.to(DEV),
                                         (1.0 / sigma.detach()).reshape(1).to(DEV))
    assert H.rel_err(out.cpu(), w.grad) < 1e-5


# ------------------------------------------------- optimizer

@pytest.mark.parametrize('betas,wd', [((0.0, 0.999), 0.0), ((0.9, 0.999), 0.0), ((0.5, 0.9), 0.01)])
def test_fused_adam_matches_torch_adam(betas, wd):
    """imx_adam_multi_tensor_f32 against torch.optim.Adam (the reference's 'adam' optimizer,
    utils/trainer.py:261-306) over three steps, odd sizes crossing chunk / vector boundaries and
    more tensors than one launch's table holds; then the state dicts are swapped."""
    from imaginaire_b200.optim import FusedAdam
    g = torch.Generator().manual_seed(51)
    sizes = [(1,), (7,), (3, 5, 5), (32768,), (32769,), (100003,), (1200, 1201)] + [(13,)] * 100
    ours = [torch.randn(*s, generator=g).to(DEV).requires_grad_(True) for s in sizes]
    ref = [p.detach().clone().requires_grad_(True) for p in ours]
    opt_o = FusedAdam(ours, lr=1e-3, betas=betas, eps=1e-8, weight_decay=wd)
    opt_r = torch.optim.Adam(ref, lr=1e-3, betas=betas, eps=1e-8, weight_decay=wd)
    for step in range(3):
        for a, b in zip(ours, ref):
            gr = torch.randn(a.shape, generator=g).to(DEV) * (10.0 ** (step - 1))
            a.grad = gr.clone()
            b.grad = gr.clone()
        opt_o.step()
        opt_r.step()
    for a, b in zip(ours, ref):
        assert H.rel_err(a.detach().cpu(), b.detach().cpu()) < 2e-6
        assert H.rel_err(opt_o.state[a]['exp_avg'].cpu(), opt_r.state[b]['exp_avg'].cpu()) < 2e-6
        assert H.rel_err(opt_o.state[a]['exp_avg_sq'].cpu(), opt_r.state[b]['exp_avg_sq'].cpu()) < 2e-6
        assert float(opt_o.state[a]['step']) == float(opt_r.state[b]['step']) == 3.0
    # checkpoints are interchangeable
    opt_r.load_state_dict(opt_o.state_dict())
    opt_o.load_state_dict(opt_r.state_dict())
    for a in ours:
        a.grad = torch.ones_like(a)
    opt_o.step()
    assert float(opt_o.state[ours[0]]['step']) == 4.0
    cpu_p = torch.zeros(3, requires_grad=True)
    cpu_p.grad = torch.ones(3)
    with pytest.raises(RuntimeError):
        FusedAdam([cpu_p]).step()       # no CPU fallback


@pytest.mark.parametrize('shape', [(2, 16, 8, 12), (1, 8, 1, 3), (3, 64, 5, 1), (2, 256, 16, 32)])
def test_bilinear_up2x_add_vs_torch(shape):
    """FPSE top-down step: bilinear x2 (align_corners=False) + lateral, forward and both gradients,
    against F.interpolate in fp32 on the same bf16-rounded inputs."""
    from imaginaire_b200 import functional as NF
    n, c, h, w = shape
    g = torch.Generator().manual_seed(81)
    src = torch.randn(n, c, h, w, generator=g).bfloat16().float()
    lat = torch.randn(n, c, 2 * h, 2 * w, generator=g).bfloat16().float()
    go = torch.randn(n, c, 2 * h, 2 * w, generator=g).bfloat16().float()
    sd, ld = nhwc(src).requires_grad_(True), nhwc(lat).requires_grad_(True)
    out = NF.bilinear_up2x_add(sd, ld)
    sr, lr = src.clone().requires_grad_(True), lat.clone().requires_grad_(True)
    ref = F.interpolate(sr, scale_factor=2, mode='bilinear', align_corners=False) + lr
    assert H.rel_err(out.float().cpu(), ref.detach()) < 4e-3
    out.backward(nhwc(go))
    ref.backward(go)
    assert H.rel_err(sd.grad.float().cpu(), sr.grad) < 4e-3
    assert torch.equal(ld.grad.float().cpu(), go)
    only_up = NF.bilinear_up2x_add(nhwc(src))
    assert H.rel_err(only_up.float().cpu(), F.interpolate(src, scale_factor=2, mode='bilinear', align_corners=False)) < 4e-3


@pytest.mark.parametrize('case', [(2, 64, 64, 16, 32, 3, 1, 1), (1, 40, 72, 13, 11, 5, 1, 2), (4, 256, 128, 32, 64, 1, 1, 0),
                                  (2, 130, 70, 9, 9, 4, 2, 1)])
def test_wgrad_with_fused_spectral_norm_inner_product(case):
    """ops.conv2d_wgrad(sn=...): <G, W> is formed while the split-K partial sums are reduced
    (imx_conv2d_igemm_wgrad_dot_bf16), then G / sigma - (<G, W> / sigma^2) u v^T.  Against the same two steps on
    the CPU in float64 from torch's weight gradient."""
    from imaginaire_b200 import ops
    n, cin, cout, h, w, k, s, p = case
    g = torch.Generator().manual_seed(141)
    ho, wo_ = (h + 2 * p - k) // s + 1, (w + 2 * p - k) // s + 1
    x = torch.randn(n, cin, h, w, generator=g).bfloat16().float()
    dy = torch.randn(n, cout, ho, wo_, generator=g).bfloat16().float()
    wt = torch.randn(cout, cin, k, k, generator=g)
    u = torch.nn.functional.normalize(torch.randn(cout, generator=g), dim=0)
    v = torch.nn.functional.normalize(torch.randn(cin * k * k, generator=g), dim=0)
    inv_sigma = torch.tensor([1.0 / 3.7])
    cin_p, cout_p = ops.round_up(cin, 8), ops.round_up(cout, 8)
    xp = torch.zeros(n, cin_p, h, w); xp[:, :cin] = x
    dp = torch.zeros(n, cout_p, ho, wo_); dp[:, :cout] = dy
    out = torch.full((cout, cin, k, k), float('nan'), device=DEV)
    got = ops.conv2d_wgrad(nhwc(xp), nhwc(dp), cin, cout, k, k, s, p, out=out,
                           sn=(wt.to(DEV), u.to(DEV), v.to(DEV), inv_sigma.to(DEV)))
    assert got.data_ptr() == out.data_ptr()
    G = torch.nn.grad.conv2d_weight(x, (cout, cin, k, k), dy, stride=s, padding=p).double()
    dot = (G * wt.double()).sum()
    ref = G * float(inv_sigma) - dot * float(inv_sigma) ** 2 * torch.outer(u.double(), v.double()).view_as(G)
    assert H.rel_err(got.cpu(), ref) < 2e-4


def test_fused_adam_device_step_counters_match_torch_adam():
    """FusedAdam(capturable=True): step numbers on the device, bias corrections formed in the kernel
    (imx_adam_multi_tensor_f32_capturable) -- the form a CUDA-graph replay needs.  Five steps against torch.optim.Adam;
    an optimizer that already stepped with host counters converts with make_capturable()."""
    from imaginaire_b200.optim import FusedAdam
    g = torch.Generator().manual_seed(71)
    shapes = [(64, 32, 3, 3), (128,), (7, 5), (40000,)]
    ps = [torch.randn(*s, generator=g) for s in shapes]
    ours = [torch.nn.Parameter(p.clone().to(DEV)) for p in ps]
    mixed = [torch.nn.Parameter(p.clone().to(DEV)) for p in ps]
    ref = [torch.nn.Parameter(p.clone().to(DEV)) for p in ps]
    kw = dict(lr=3e-3, betas=(0.5, 0.99), eps=1e-8)
    o1, o2, o3 = FusedAdam(ours, capturable=True, **kw), FusedAdam(mixed, **kw), torch.optim.Adam(ref, **kw)
    for it in range(5):
        if it == 2:
            o2.make_capturable()
        for a, m, r in zip(ours, mixed, ref):
            grad = torch.randn(a.shape, generator=g).to(DEV)
            a.grad, m.grad, r.grad = grad.clone(), grad.clone(), grad.clone()
        o1.step(); o2.step(); o3.step()
    for a, m, r in zip(ours, mixed, ref):
        assert H.rel_err(a.detach().cpu(), r.detach().cpu()) < 2e-6
        assert H.rel_err(m.detach().cpu(), r.detach().cpu()) < 2e-6
    st = o1.state[ours[0]]['step']
    assert st.is_cuda and float(st) == 5.0 and float(o2.state[mixed[0]]['step']) == 5.0


# ------------------------------------------------- FlowNet2 operators

@pytest.mark.parametrize('shape', [(2, 3, 17, 23), (1, 2, 64, 96), (3, 64, 5, 7), (1, 1, 1, 1)])
def test_channelnorm_bit_exact_vs_oracle(shape):
    from oracle import flownet_ops_oracle as fo
    from imaginaire_b200.third_party.channelnorm import ChannelNorm
    g = torch.Generator().manual_seed(61)
    x = torch.randn(*shape, generator=g)
    xd = x.to(DEV).requires_grad_(True)
    y = ChannelNorm()(xd)
    ref = fo.channelnorm_forward(x.numpy())
    assert np.array_equal(y.detach().cpu().numpy(), ref)
    go = torch.randn(y.shape, generator=g)
    y.backward(go.to(DEV))
    assert np.array_equal(xd.grad.cpu().numpy(), fo.channelnorm_backward(x.numpy(), ref, go.numpy()))
    with pytest.raises(Exception):
        ChannelNorm(norm_deg=1)(xd)


@pytest.mark.parametrize('case', [(2, 64, 12, 16, 20, 1, 20, 1, 2),      # FlowNetC geometry
                                  (1, 40, 9, 11, 4, 1, 4, 1, 2),          # channels not a multiple of 32
                                  (2, 32, 10, 14, 4, 1, 4, 1, 1),
                                  (1, 16, 13, 15, 7, 3, 5, 1, 2)])        # 3x3 patches
def test_correlation_bit_exact_vs_oracle(case):
    """Forward and both gradients of the FlowNetC correlation.  The kernels keep the reference's
    lane assignment and reduction order, the oracle emulates the 32 lanes: bit-exact."""
    from oracle import flownet_ops_oracle as fo
    from imaginaire_b200.third_party.correlation import Correlation
    b, c, h, w, pad, k, md, s1, s2 = case
    g = torch.Generator().manual_seed(62)
    x1, x2 = torch.randn(b, c, h, w, generator=g), torch.randn(b, c, h, w, generator=g)
    d1, d2 = x1.to(DEV).requires_grad_(True), x2.to(DEV).requires_grad_(True)
    out = Correlation(pad_size=pad, kernel_size=k, max_displacement=md, stride1=s1, stride2=s2)(d1, d2)
    ref = fo.correlation_forward(x1.numpy(), x2.numpy(), pad, k, md, s1, s2)
    assert tuple(out.shape) == ref.shape
    assert np.array_equal(out.detach().cpu().numpy(), ref)
    go = torch.randn(out.shape, generator=g)
    out.backward(go.to(DEV))
    g1, g2 = fo.correlation_backward(x1.numpy(), x2.numpy(), go.numpy(), pad, k, md, s1, s2)
    assert np.array_equal(d1.grad.cpu().numpy(), g1)
    assert np.array_equal(d2.grad.cpu().numpy(), g2)


def test_correlation_rejects_geometry_the_reference_reads_out_of_bounds_for():
    from imaginaire_b200 import ops, _C
    x = torch.randn(1, 8, 8, 8, device=DEV)
    with pytest.raises(_C.ImxError):
        ops.correlation_fwd(x, x, 4, 3, 4, 1, 2)          # 3x3 patches need pad >= max_displacement + 1
    out = ops.correlation_fwd(x, x, 4, 1, 4, 2, 2)        # stride1 = 2 forward is fine ...
    with pytest.raises(_C.ImxError):
        ops.correlation_bwd(x, x, torch.zeros_like(out), 4, 1, 4, 2, 2)   # ... its gradient is not
    with pytest.raises(RuntimeError):
        ops.correlation_fwd(x.cpu(), x.cpu(), 4, 1, 4, 1, 2)              # no CPU fallback


@pytest.mark.parametrize('world', [2, 3, 8])
def test_peer_reduce_scatter_and_all_gather_average_the_arenas(world):
    """csrc/peer_grads.cu with the ranks' arenas emulated by tensors on one GPU: after the reduce-scatter of every
    rank and then the all-gather of every rank, every arena holds the average of the original arenas on the listed
    segments (ragged sizes: chunk boundaries are rounded to 4 floats) and is untouched elsewhere."""
    import ctypes
    from imaginaire_b200 import _C, ops
    g = torch.Generator().manual_seed(91)
    sizes = [4096, 7, 130, 100000, 1, 64 * 9 * 33]
    offsets, top = [], 64
    for n in sizes:
        offsets.append(top)
        top += (n + 63) // 64 * 64 + 64                 # a gap after every segment
    arenas = [torch.randn(top, generator=g).to(DEV) for _ in range(world)]
    orig = [a.clone() for a in arenas]
    segs = torch.tensor([[o, n] for o, n in zip(offsets, sizes)], dtype=torch.int64, device=DEV)
    table = (ctypes.c_void_p * world)(*[a.data_ptr() for a in arenas])
    lib = _C.lib()
    for r in range(world):
        lib.call('imx_peer_reduce_scatter_f32', table, world, r, segs.data_ptr(), len(sizes), max(sizes), 1.0 / world, ops._stream())
    for r in range(world):
        lib.call('imx_peer_all_gather_f32', table, world, r, segs.data_ptr(), len(sizes), max(sizes), ops._stream())
    torch.cuda.synchronize()
    mean = torch.stack(orig).mean(0)
    mask = torch.zeros(top, dtype=torch.bool, device=DEV)
    for o, n in zip(offsets, sizes):
        mask[o:o + n] = True
    for r in range(world):
        assert H.rel_err(arenas[r][mask].cpu(), mean[mask].cpu()) < 1e-6
        assert torch.equal(arenas[r][~mask], orig[r][~mask])
        assert torch.equal(arenas[r][mask], arenas[0][mask])         # bit-identical on every rank


# ------------------------------------------------- SyncBatchNorm exchange over peer memory

@pytest.mark.parametrize('world', [2, 4, 8])
def test_peer_exchange_allreduce_and_bn_finalize_emulated_on_one_gpu(world):
    """csrc/peer_exchange.cu with `world` ranks emulated on one GPU: one region, one stream and one single-CTA kernel
    per rank, all resident at once and waiting for each other's flags exactly as across NVLink.  Six calls in a row
    exercise both slot sets; results must be identical on every rank and equal to the sum / to bn_finalize of the
    summed statistics (torch.nn.SyncBatchNorm forward: nn/modules/_functions.py:49-86)."""
    from imaginaire_b200 import ops, _C
    from imaginaire_b200.peer_exchange import PeerExchange
    nbytes = int(_C.lib().raw('imx_peer_exchange_region_bytes')())
    regions = [torch.zeros(nbytes // 4, dtype=torch.float32, device=DEV) for _ in range(world)]
    ranks = [PeerExchange(regions=regions, rank=r) for r in range(world)]
    streams = [torch.cuda.Stream() for _ in range(world)]
    g = torch.Generator().manual_seed(81)
    torch.cuda.synchronize()
    for call in range(6):
        c = [64, 2048, 40, 4096, 8, 1024][call]
        if call % 2 == 0:
            vecs = [torch.randn(2, 1, c, generator=g).to(DEV) for _ in range(world)]
            torch.cuda.synchronize()
            outs = []
            for r in range(world):
                with torch.cuda.stream(streams[r]):
                    outs.append(ranks[r].allreduce_sum(vecs[r]))
            torch.cuda.synchronize()
            expect = torch.stack(vecs).sum(0)
            for o in outs:
                assert torch.equal(o, outs[0])                       # same summation order everywhere
                assert H.rel_err(o.cpu(), expect.cpu()) < 1e-6
        else:
            xs = [torch.randn(3, c, 5, 7, generator=g) * 2 + 0.5 for _ in range(world)]
            sums = [torch.stack([x.sum((0, 2, 3)), (x * x).sum((0, 2, 3))]).view(2, 1, c).to(DEV) for x in xs]
            count = world * 3 * 5 * 7
            rm = [torch.zeros(c, device=DEV) for _ in range(world)]
            rv = [torch.ones(c, device=DEV) for _ in range(world)]
            torch.cuda.synchronize()
            res = []
            for r in range(world):
                with torch.cuda.stream(streams[r]):
                    res.append(ranks[r].bn_finalize(sums[r], count, 1e-5, rm[r], rv[r], 0.1))
            torch.cuda.synchronize()
            allx = torch.cat(xs, 0)
            mean = allx.mean((0, 2, 3))
            var = allx.var((0, 2, 3), unbiased=False)
            for r in range(world):
                assert torch.equal(res[r][0], res[0][0]) and torch.equal(res[r][1], res[0][1])
                assert H.rel_err(res[r][0].cpu(), mean) < 1e-5
                assert H.rel_err(res[r][1].cpu(), torch.rsqrt(var + 1e-5)) < 1e-4
                assert H.rel_err(rm[r].cpu(), 0.1 * mean) < 1e-5
                assert H.rel_err(rv[r].cpu(), 0.9 + 0.1 * allx.var((0, 2, 3), unbiased=True)) < 1e-4
    with pytest.raises(_C.ImxError, match='exceed'):
        ranks[0].allreduce_sum(torch.zeros(8193, device=DEV))


# ------------------------------------------------- data-parallel weight gradient over peer memory

@pytest.mark.parametrize('case', [(2, 64, 96, 16, 24, 3, 1, 1, 2), (1, 40, 72, 13, 11, 5, 1, 2, 3), (2, 128, 256, 8, 8, 1, 1, 0, 8)])
def test_wgrad_peer_reduce_and_gather_equal_the_averaged_gradient(case):
    """imx_conv2d_igemm_wgrad_partials_bf16 + imx_wgrad_reduce_peers_f32 + imx_gather_rows_peers_f32
    (grad_sync.py's data-parallel path) with the 'ranks' emulated by separate buffers on one GPU:
    every rank must end up with the mean of the per-rank gradients of the ordinary entry point."""
    import ctypes
    from imaginaire_b200 import ops, _C
    n, cin, cout, h, w, k, s, p, world = case
    lib = _C.lib()
    g = torch.Generator().manual_seed(71)
    cin_p, cout_p = ops.round_up(cin, 8), ops.round_up(cout, 8)
    ho, wo_ = (h + 2 * p - k) // s + 1, (w + 2 * p - k) // s + 1
    xs = [nhwc(torch.randn(n, cin_p, h, w, generator=g)) for _ in range(world)]
    dys = [nhwc(torch.randn(n, cout_p, ho, wo_, generator=g)) for _ in range(world)]
    ref = sum(ops.conv2d_wgrad(x, dy, cin, cout, k, k, s, p) for x, dy in zip(xs, dys)) / world
    nbytes = lib.raw('imx_conv2d_wgrad_workspace_bytes')(n, h, w, cin, cin_p, cout, cout_p, k, k, s, p)
    wss = [torch.empty(int(nbytes), dtype=torch.uint8, device=DEV) for _ in range(world)]
    stream = torch.cuda.current_stream().cuda_stream
    plan = (ctypes.c_int * 3)()
    for x, dy, ws in zip(xs, dys, wss):
        lib.call('imx_conv2d_igemm_wgrad_partials_bf16', x.data_ptr(), dy.data_ptr(), ws.data_ptr(), int(nbytes), plan,
                 n, h, w, cin, cin_p, cout, cout_p, k, k, s, p, stream)
    numel = cout * cin * k * k
    slot = 64                                      # the tensor sits at an offset inside each 'arena'
    arenas = [torch.full((slot + numel,), float('nan'), device=DEV) for _ in range(world)]
    peers = (ctypes.c_void_p * world)(*[ws.data_ptr() for ws in wss])
    for r in range(world):
        lib.call('imx_wgrad_reduce_peers_f32', peers, world, arenas[r].data_ptr() + 4 * slot, r * cout // world,
                 (r + 1) * cout // world, cout, cin, k * k, int(plan[0]), int(plan[1]), int(plan[2]), 1.0 / world, 0, stream)
    descs = torch.tensor([[slot, cout, cin * k * k]], dtype=torch.int64).to(DEV)
    apeers = (ctypes.c_void_p * world)(*[a.data_ptr() for a in arenas])
    for r in range(world):
        lib.call('imx_gather_rows_peers_f32', apeers, world, r, descs.data_ptr(), 1, numel, stream)
    torch.cuda.synchronize()
    for r in range(world):
        got = arenas[r][slot:].view(cout, cin, k, k)
        assert bool(torch.isfinite(got).all())
        assert H.rel_err(got.cpu(), ref.cpu()) < 2e-6, r
        assert torch.equal(arenas[r][slot:], arenas[0][slot:])          # every rank holds the same bits


@pytest.mark.parametrize('shape,nhwc', [((2, 16, 8, 12), True), ((1, 8, 1, 3), True), ((3, 7, 5, 3), False),
                                        ((8, 64, 64, 96), True), ((1, 1, 1, 1), False)])
def test_l1_loss_mean_vs_torch(shape, nhwc):
    """Fused feature-map L1 (value and gradient) against the torch expression it replaces, in fp32."""
    from imaginaire_b200 import functional as NF
    torch.manual_seed(11)
    dev = torch.device('cuda:0')
    a = torch.randn(*shape, device=dev).bfloat16()
    b = torch.randn(*shape, device=dev).bfloat16()
    b.view(-1)[::3] = a.view(-1)[::3]                     # exact ties: sign(0) = 0
    if nhwc:
        a = a.contiguous(memory_format=torch.channels_last)
        b = b.contiguous(memory_format=torch.channels_last)
    a1 = a.clone().requires_grad_(True)
    loss = NF.l1_loss_mean(a1, b) * 3.0
    loss.backward()
    a2 = a.clone().requires_grad_(True)
    ref = torch.nn.functional.l1_loss(a2.float(), b.float()) * 3.0
    ref.backward()
    assert loss.dtype == torch.float32 and loss.dim() == 0
    # fp32 sums in two different orders (atomics across CTAs here, ATen's tree there)
    assert abs(float(loss) - float(ref)) <= 2e-5 * max(1.0, abs(float(ref)))
    assert a1.grad.dtype == torch.bfloat16 and a1.grad.stride() == a.stride()
    assert torch.equal(a1.grad, a2.grad)                  # both are sign * (3 / n) rounded to bf16 once
    # gradient with respect to the second argument
    b3 = b.clone().requires_grad_(True)
    NF.l1_loss_mean(a, b3).backward()
    b4 = b.clone().requires_grad_(True)
    torch.nn.functional.l1_loss(a.float(), b4.float()).backward()
    assert torch.equal(b3.grad, b4.grad)


def test_l1_loss_mean_other_inputs_take_the_torch_expression():
    from imaginaire_b200 import functional as NF
    dev = torch.device('cuda:0')
    a = torch.randn(2, 3, 4, 5, device=dev)
    b = torch.randn(2, 3, 4, 5, device=dev)
    assert torch.equal(NF.l1_loss_mean(a, b), torch.nn.functional.l1_loss(a, b))
    ab = a.bfloat16()
    bb = b.bfloat16().contiguous(memory_format=torch.channels_last)      # layouts differ
    assert torch.equal(NF.l1_loss_mean(ab, bb), torch.nn.functional.l1_loss(ab.float(), bb.float()))


@pytest.mark.parametrize('shape', [(2, 16, 8, 12), (1, 8, 2, 2), (3, 64, 7, 9), (2, 128, 32, 48), (1, 8, 3, 5)])
def test_max_pool2x2_vs_torch(shape):
    """Native NHWC bf16 2x2 max pooling: values and gradient routing (ties -> first maximum) as ATen."""
    from imaginaire_b200 import functional as NF
    torch.manual_seed(5)
    dev = torch.device('cuda:0')
    x = torch.randn(*shape, device=dev).relu_().bfloat16().contiguous(memory_format=torch.channels_last)   # ReLU: many tied zeros
    x1 = x.clone().requires_grad_(True)
    y = NF.max_pool2x2(x1)
    x2 = x.clone().requires_grad_(True)
    ref = torch.nn.functional.max_pool2d(x2, kernel_size=2, stride=2)
    assert y.shape == ref.shape and torch.equal(y, ref)
    if y.numel():
        g = torch.randn_like(ref)
        y.backward(g)
        ref.backward(g)
        assert torch.equal(x1.grad, x2.grad)


@pytest.mark.parametrize('shape', [(2, 5, 16, 24), (1, 38, 64, 96), (3, 1, 7, 9), (2, 3, 2, 2), (1, 2, 3, 2)])
def test_bilinear_half_align_corners_vs_torch(shape):
    """The discriminator pyramid's 0.5x bilinear resize (align_corners=True) against F.interpolate."""
    from imaginaire_b200 import functional as NF
    torch.manual_seed(9)
    dev = torch.device('cuda:0')
    x = torch.randn(*shape, device=dev)
    x1 = x.clone().requires_grad_(True)
    y = NF.bilinear_half_align_corners(x1)
    x2 = x.clone().requires_grad_(True)
    ref = torch.nn.functional.interpolate(x2, scale_factor=0.5, mode='bilinear', align_corners=True)
    assert y.shape == ref.shape
    assert torch.allclose(y, ref, rtol=1e-6, atol=1e-6)          # same formula; only FMA contraction may differ
    g = torch.randn_like(ref)
    y.backward(g)
    ref.backward(g)
    assert torch.allclose(x1.grad, x2.grad, rtol=1e-6, atol=1e-6)


@pytest.mark.parametrize('shape', [(2, 16, 8, 12), (1, 8, 1, 3), (3, 256, 5, 7), (2, 264, 4, 4)])
def test_channel_dot_vs_torch(shape):
    """FPSE projection term: per-pixel dot over channels and its gradients against torch in fp32."""
    from imaginaire_b200 import functional as NF
    torch.manual_seed(13)
    dev = torch.device('cuda:0')
    a = torch.randn(*shape, device=dev).bfloat16().contiguous(memory_format=torch.channels_last)
    b = torch.randn(*shape, device=dev).bfloat16().contiguous(memory_format=torch.channels_last)
    a1, b1 = a.clone().requires_grad_(True), b.clone().requires_grad_(True)
    y = NF.channel_dot(a1, b1)
    a2, b2 = a.float().requires_grad_(True), b.float().requires_grad_(True)
    ref = torch.mul(a2, b2).sum(dim=1, keepdim=True)
    assert y.shape == ref.shape and y.dtype == torch.bfloat16
    assert H.rel_err(y.float(), ref) < 1e-2                      # one bf16 rounding of the fp32 sum
    g = torch.randn_like(ref).bfloat16()
    y.backward(g)
    ref.backward(g.float())
    assert H.rel_err(a1.grad.float(), a2.grad) < 1e-2 and H.rel_err(b1.grad.float(), b2.grad) < 1e-2



# ---------------------------------------------------------------- filter packs (layout.cu: pack_weight_rows_kernel)
PACK_CASES = [  # cout, cin, k, padded inner (None: next multiple of 8)
    (64, 3, 3, None), (40, 35, 4, 40), (128, 38, 4, None), (3, 256, 5, None), (130, 72, 1, None), (257, 129, 3, 136),
    (96, 64, 7, None), (512, 128, 5, None), (1, 1, 3, None), (300, 520, 3, None), (33, 17, 2, 24), (256, 384, 3, None),
    (128, 256, 4, None), (264, 136, 4, None), (130, 256, 5, 256),
]


@pytest.mark.gpu
@pytest.mark.parametrize('cout,cin,k,inner_p', PACK_CASES)
def test_filter_packs_are_bit_exact(cout, cin, k, inner_p):
    """packed[co][tap][ci_p] = bf16(w * s) and the flipped / transposed dgrad filter, against torch indexing:
    ragged tiles (inner % 128 != 0), fewer rows than a CTA's row group, even and odd tap counts, padding columns."""
    from imaginaire_b200 import ops
    g = torch.Generator().manual_seed(cout * 1000 + cin * 10 + k)
    w = torch.randn(cout, cin, k, k, generator=g).to(DEV)
    s = torch.tensor([0.731], device=DEV)
    for flip in (False, True):
        inner = cout if flip else cin
        ip = inner_p if (inner_p is not None and inner_p >= inner) else (inner + 7) // 8 * 8
        got = ops.pack_conv_weight(w, inv_sigma=s, cin_p=ip, transpose_flip=flip)
        ws = (w * s).bfloat16()
        ref = ws.flip(2, 3).permute(1, 2, 3, 0) if flip else ws.permute(0, 2, 3, 1)
        want = torch.zeros(got.shape, dtype=torch.bfloat16, device=DEV)
        want[..., :inner] = ref
        assert torch.equal(got.view(torch.int16), want.view(torch.int16)), 'flip=%s' % flip


@pytest.mark.gpu
@pytest.mark.parametrize('c,ch,k', [(64, 128, 5), (96, 128, 3), (200, 40, 5), (32, 16, 1), (512, 128, 5)])
def test_interleaved_gamma_beta_pack_is_bit_exact(c, ch, k):
    from imaginaire_b200 import ops
    g = torch.Generator().manual_seed(c + ch + k)
    wg, wb = torch.randn(c, ch, k, k, generator=g).to(DEV), torch.randn(c, ch, k, k, generator=g).to(DEV)
    sg, sb = torch.tensor([1.3], device=DEV), torch.tensor([0.4], device=DEV)
    got = ops.pack_spade_gb_weight(wg, wb, sg, sb)
    cr = (c + 63) // 64 * 64
    want = torch.zeros(2 * cr, k, k, (ch + 7) // 8 * 8, dtype=torch.bfloat16, device=DEV)
    cs = torch.arange(c, device=DEV)
    rows = (cs // 64) * 128 + cs % 64
    want[rows, :, :, :ch] = (wg * sg).bfloat16().permute(0, 2, 3, 1)
    want[rows + 64, :, :, :ch] = (wb * sb).bfloat16().permute(0, 2, 3, 1)
    assert torch.equal(got.view(torch.int16), want.view(torch.int16))


@pytest.mark.gpu
@pytest.mark.parametrize('shape,cp', [((2, 38, 37, 53), 40), ((8, 3, 64, 128), 8), ((1, 35, 1, 5), 40), ((3, 64, 16, 24), 64),
                                      ((2, 9, 128, 1), 16), ((2, 72, 19, 21), 72), ((1, 130, 8, 40), 136)])
def test_nchw_to_nhwc_conversion_is_bit_exact(shape, cp):
    """fp32 NCHW -> bf16 NHWC with zero-filled padding channels: the 128-pixel strip kernel (Cp <= 64) with ragged
    strips, and the 32 x 32 tile kernel for wider tensors."""
    from imaginaire_b200 import ops
    g = torch.Generator().manual_seed(sum(shape))
    x = torch.randn(*shape, generator=g).to(DEV)
    got = ops.nchw_f32_to_nhwc_bf16(x, cp=cp)
    assert got.shape == (shape[0], cp, shape[2], shape[3]) and got.is_contiguous(memory_format=torch.channels_last)
    want = torch.zeros(shape[0], cp, shape[2], shape[3], dtype=torch.bfloat16, device=DEV)
    want[:, :shape[1]] = x.bfloat16()
    assert torch.equal(got.contiguous().view(torch.int16), want.view(torch.int16))


@pytest.mark.gpu
@pytest.mark.parametrize('r,c', [(64, 576), (37, 100), (2048, 1152), (5, 27), (1, 4)])
def test_spectral_norm_weight_gradient_correction(r, c):
    """G <- G / sigma - (<G, W> / sigma^2) u v^T (imx_spectral_norm_weight_grad_f32): the row-walking float4 kernel
    (C % 4 == 0) and the scalar one, against float64."""
    from imaginaire_b200 import ops
    g = torch.Generator().manual_seed(r + c)
    G, W = torch.randn(r, c, generator=g), torch.randn(r, c, generator=g)
    u = torch.nn.functional.normalize(torch.randn(r, generator=g), dim=0)
    v = torch.nn.functional.normalize(torch.randn(c, generator=g), dim=0)
    inv_sigma = torch.tensor([0.37])
    Gd = G.clone().to(DEV)
    vb = torch.zeros(c + 1, device=DEV)          # v at an odd offset of a larger buffer, as the batched power iteration leaves it
    vb[1:] = v.to(DEV)
    ops.spectral_norm_weight_grad_(Gd, W.to(DEV), u.to(DEV), vb[1:], inv_sigma.to(DEV))
    dot = (G.double() * W.double()).sum()
    ref = G.double() * 0.37 - dot * 0.37 ** 2 * torch.outer(u.double(), v.double())
    assert H.rel_err(Gd.cpu(), ref) < 1e-5
