"""BASELINE.json configs[1]: SPADE ResBlk microbench sweep, plus the HBM-bound kernels
(statistics, fused modulate, flow warps) on their own, on one B200.

For each (channels, resolution) a `Res2dBlock` with the C3 SPADE parameters (hidden 128, k5,
separate projection, sync batch norm, spectral norm) runs forward under no_grad (fused
gamma/beta path) and forward+backward; beside it the same block written with plain torch
ops (cuDNN bf16 channels_last convs + ATen batch norm / elementwise) as the on-GPU reference.
Prints one JSON line per case and a final summary line; CUDA-event timing, 3 warm-ups.

    python tools/resblk_sweep.py [--quick | --full]      (--full: SURVEY 8(d)'s grid C in 64..1024 x 16x32..256x512)
"""
import json
import os
import sys
import types

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F

from imaginaire_b200 import functional as NF
from imaginaire_b200 import ops
from imaginaire_b200.layers import Res2dBlock

PEAK_HBM, PEAK_TF = 6549.4, 1633.4
try:
    with open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'MEASURED_PEAKS.json')) as f:
        _p = json.load(f)
        PEAK_HBM, PEAK_TF = _p['hbm_gbs'], _p['bf16_tflops']
except OSError:
    pass


def timeit(fn, iters=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters):
        fn()
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) / iters


def nhwc(t):
    return t.bfloat16().contiguous(memory_format=torch.channels_last)


def make_block(cin, cout, labels=35):
    an = types.SimpleNamespace(num_filters=128, kernel_size=5, separate_projection=True, cond_dims=labels,
                               weight_norm_type='spectral', activation_norm_type='sync_batch',
                               activation_norm_params=types.SimpleNamespace(affine=True))
    blk = Res2dBlock(cin, cout, kernel_size=3, padding=1, bias=[True, True, False], weight_norm_type='spectral',
                     activation_norm_type='spatially_adaptive', activation_norm_params=an,
                     skip_activation_norm=True, nonlinearity='leakyrelu', order='NACNAC')
    return blk.cuda().train()


def block_flops(cin, cout, n, h, w, labels=35, hidden=128):
    px = n * h * w
    mid = min(cin, cout)
    main = 2 * px * 9 * (cin * mid + mid * cout) + (2 * px * cin * cout if cin != cout else 0)
    norms = [cin, mid] + ([cin] if cin != cout else [])
    spade = sum(2 * px * 25 * (labels * hidden + 2 * hidden * c) for c in norms)
    return main + spade


def torch_block(blk, x, label):
    """The same block with plain torch ops on bf16 channels_last tensors (cuDNN / ATen)."""
    def sn_w(conv):
        return conv.weight_orig / torch.dot(conv.weight_u, torch.mv(conv.weight_orig.flatten(1), conv.weight_v))

    def spade(norm, t):
        lab = F.interpolate(label, size=t.shape[2:], mode='nearest').bfloat16()
        hid = F.relu(F.conv2d(lab, sn_w(norm.mlps[0][0].layers['conv']).bfloat16(), norm.mlps[0][0].layers['conv'].bias.bfloat16(), padding=2))
        g = F.conv2d(hid, sn_w(norm.gammas[0].layers['conv']).bfloat16(), norm.gammas[0].layers['conv'].bias.bfloat16(), padding=2)
        b = F.conv2d(hid, sn_w(norm.betas[0].layers['conv']).bfloat16(), norm.betas[0].layers['conv'].bias.bfloat16(), padding=2)
        tn = F.batch_norm(t, None, None, norm.norm.weight.bfloat16(), norm.norm.bias.bfloat16(), True)
        return tn * (1 + g) + b

    def cb(block, t, act=True):
        t = spade(block.layers['norm'], t)
        if act:
            t = F.leaky_relu(t, 0.2)
        c = block.layers['conv']
        return F.conv2d(t, sn_w(c).bfloat16(), None if c.bias is None else c.bias.bfloat16(), padding=c.padding)

    dx = cb(blk.conv_block_1, cb(blk.conv_block_0, x))
    xs = cb(blk.conv_block_s, x, act=False) if blk.learn_shortcut else x
    return xs + dx


def main():
    quick = '--quick' in sys.argv
    torch.manual_seed(0)
    n = 8
    label = torch.zeros(n, 35, 256, 512, device='cuda').scatter_(
        1, torch.randint(0, 35, (n, 1, 32, 64), device='cuda').repeat_interleave(8, 2).repeat_interleave(8, 3), 1.0)
    cases = [(1024, 16, 32), (1024, 32, 64), (512, 64, 128), (512, 128, 256), (256, 128, 256), (128, 256, 512),
             (64, 256, 512)]
    if quick:
        cases = cases[1:4]
    if '--full' in sys.argv:
        cases = [(c, h, w) for c in (64, 128, 256, 512, 1024) for h, w in ((16, 32), (32, 64), (64, 128), (128, 256), (256, 512))]
    results = []
    for c, h, w in cases:
        for cout in ((c, c // 2) if '--full' not in sys.argv else (c,)):
            blk = make_block(c, cout)
            x = nhwc((torch.randn(n, c, h, w, device='cuda') * 1.5 + 0.3))
            fl = block_flops(c, cout, n, h, w)
            with torch.no_grad():
                t_inf = timeit(lambda: blk(x, label))
                t_ref = timeit(lambda: torch_block(blk, x, label))
            xg = x.clone().requires_grad_(True)

            def train():
                y = blk(xg, label)
                y.backward(y)
            t_train = timeit(train, iters=5)
            xr = x.clone().requires_grad_(True)

            def train_ref():
                y = torch_block(blk, xr, label)
                y.backward(y)
            t_train_ref = timeit(train_ref, iters=5)
            r = {'case': 'Res2dBlock %d->%d @%dx%d N=%d' % (c, cout, h, w, n), 'fwd_ms': t_inf, 'fwd_tflops': fl / t_inf / 1e9,
                 'fwd_frac_of_peak': fl / t_inf / 1e9 / PEAK_TF, 'torch_cudnn_fwd_ms': t_ref, 'fwd_speedup_vs_torch': t_ref / t_inf,
                 'train_ms': t_train, 'train_tflops': 3 * fl / t_train / 1e9, 'torch_cudnn_train_ms': t_train_ref,
                 'train_speedup_vs_torch': t_train_ref / t_train}
            print(json.dumps(r), flush=True)
            results.append(r)
            del blk
            torch.cuda.empty_cache()
    # HBM-bound kernels on their own (SURVEY 8(d) accounting)
    for c, h, w in [(512, 128, 256), (1024, 64, 128), (256, 256, 512)]:
        x = nhwc(torch.randn(n, c, h, w, device='cuda'))
        g, b = nhwc(torch.randn(n, c, h, w, device='cuda')), nhwc(torch.randn(n, c, h, w, device='cuda'))
        nb = x.numel() * 2
        t = timeit(lambda: ops.bn_stats(x, False))
        sums = ops.bn_stats(x, False)
        mean, rstd = ops.bn_finalize(sums, n * h * w, 1e-5)
        t2 = timeit(lambda: ops.spade_modulate_fwd(x, mean, rstd, None, None, g, b, False, 1))
        r = {'case': 'norm kernels C=%d @%dx%d N=%d' % (c, h, w, n), 'bn_stats_ms': t, 'bn_stats_gbs': nb / t / 1e6,
             'bn_stats_frac': nb / t / 1e6 / PEAK_HBM, 'modulate_ms': t2, 'modulate_gbs_4tensors': 4 * nb / t2 / 1e6,
             'modulate_frac': 4 * nb / t2 / 1e6 / PEAK_HBM,
             'stats_plus_modulate_gbs_5x_accounting': 5 * nb / (t + t2) / 1e6}
        print(json.dumps(r), flush=True)
        results.append(r)
    for bsz, c in [(1, 3), (8, 3), (8, 64), (32, 4)]:
        img = torch.randn(bsz, c, 512, 1024, device='cuda')
        flow = F.interpolate(torch.randn(bsz, 2, 32, 64, device='cuda') * 8, size=(512, 1024), mode='bilinear').contiguous()
        by = (2 * c + 2) * bsz * 512 * 1024 * 4
        t = timeit(lambda: ops.resample2d_fwd(img, flow))
        t2 = timeit(lambda: ops.gridsample_warp_fwd(img, flow))
        gout = torch.randn_like(img)
        t3 = timeit(lambda: ops.gridsample_warp_bwd(img, flow, gout))
        grid = None

        def torch_resample():
            gx = torch.linspace(-1, 1, 1024, device='cuda').view(1, 1, 1, 1024).expand(bsz, 1, 512, 1024)
            gy = torch.linspace(-1, 1, 512, device='cuda').view(1, 1, 512, 1).expand(bsz, 1, 512, 1024)
            fl = torch.cat([flow[:, 0:1] / (1023 / 2.0), flow[:, 1:2] / (511 / 2.0)], 1)
            return F.grid_sample(img, (torch.cat([gx, gy], 1) + fl).permute(0, 2, 3, 1), mode='bilinear',
                                 padding_mode='border', align_corners=True)
        t4 = timeit(torch_resample)
        r = {'case': 'warp B=%d C=%d 512x1024' % (bsz, c), 'resample2d_fwd_ms': t, 'resample2d_gbs': by / t / 1e6,
             'resample2d_frac': by / t / 1e6 / PEAK_HBM, 'gridsample_warp_fwd_ms': t2, 'gridsample_warp_gbs': by / t2 / 1e6,
             'gridsample_warp_frac': by / t2 / 1e6 / PEAK_HBM, 'gridsample_warp_bwd_ms': t3,
             'bwd_gbs': (3 * c + 4) * bsz * 512 * 1024 * 4 / t3 / 1e6, 'torch_grid_sample_ms': t4, 'speedup_vs_torch': t4 / t2}
        print(json.dumps(r), flush=True)
        results.append(r)


if __name__ == '__main__':
    main()
