"""Short single-kernel workloads for `ncu --set full` captures (developer tool).

    python tools/ncu_kernels.py conv     # 3x3 512->512 conv at 8 x 128 x 256 (fprop, wgrad, fused SPADE gb)
    python tools/ncu_kernels.py hbm      # stats / modulate / warps at HBM-bound sizes
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from imaginaire_b200 import ops


def nhwc(t):
    return t.bfloat16().contiguous(memory_format=torch.channels_last)


def conv():
    torch.manual_seed(0)
    n, c, h, w = 8, 512, 128, 256
    x = nhwc(torch.randn(n, c, h, w, device='cuda'))
    wpk = ops.pack_conv_weight(torch.randn(c, c, 3, 3, device='cuda') * 0.01)
    dy = nhwc(torch.randn(n, c, h, w, device='cuda'))
    st = torch.zeros(2, 1, c, device='cuda')
    for _ in range(3):
        ops.conv2d_fprop(x, wpk, c, 3, 3, 1, 1, stats=st)
        ops.conv2d_wgrad(x, dy, c, c, 3, 3, 1, 1)
    hid = nhwc(torch.relu(torch.randn(n, 128, h, w, device='cuda')))
    wgb = ops.pack_spade_gb_weight(torch.randn(c, 128, 5, 5, device='cuda') * 0.01, torch.randn(c, 128, 5, 5, device='cuda') * 0.01)
    sums = ops.bn_stats(x, False)
    mean, rstd = ops.bn_finalize(sums, n * h * w, 1e-5)
    for _ in range(3):
        ops.spade_gb_conv_modulate(hid, wgb, None, None, x, mean, rstd, None, None, c, 5, 5, 2, False, 1)
    torch.cuda.synchronize()


def hbm():
    torch.manual_seed(0)
    n, c, h, w = 8, 512, 128, 256
    x = nhwc(torch.randn(n, c, h, w, device='cuda'))
    g = nhwc(torch.randn(n, c, h, w, device='cuda'))
    b = nhwc(torch.randn(n, c, h, w, device='cuda'))
    for _ in range(3):
        sums = ops.bn_stats(x, False)
        mean, rstd = ops.bn_finalize(sums, n * h * w, 1e-5)
        y = ops.spade_modulate_fwd(x, mean, rstd, None, None, g, b, False, 1)
        ops.spade_modulate_bwd(x, y, mean, rstd, None, None, g, b, False, 1, n * h * w)
        ops.act_bwd_bias(g, y, 1, True)
    img = torch.randn(8, 64, 512, 1024, device='cuda')
    # SURVEY 8(d) config 4 flow: bilinear upsample of N(0, 8 px) drawn at 32 x 64
    flow = torch.nn.functional.interpolate(torch.randn(8, 2, 32, 64, device='cuda') * 8.0, size=(512, 1024),
                                           mode='bilinear', align_corners=False).contiguous()
    go = torch.randn(8, 64, 512, 1024, device='cuda')
    for _ in range(2):
        ops.resample2d_fwd(img, flow)
        ops.gridsample_warp_fwd(img, flow)
        ops.resample2d_bwd(img, flow, go)
    torch.cuda.synchronize()


def warp():
    """The flow warps alone at B=8, C=64, 512x1024 on the SURVEY flow."""
    torch.manual_seed(0)
    img = torch.randn(8, 64, 512, 1024, device='cuda')
    flow = torch.nn.functional.interpolate(torch.randn(8, 2, 32, 64, device='cuda') * 8.0, size=(512, 1024),
                                           mode='bilinear', align_corners=False).contiguous()
    go = torch.randn(8, 64, 512, 1024, device='cuda')
    for _ in range(2):
        ops.resample2d_fwd(img, flow)
        ops.gridsample_warp_fwd(img, flow)
        ops.resample2d_bwd(img, flow, go)
    torch.cuda.synchronize()


def epi():
    """Epilogue-dominated convs: 1x1 256->512, 5x5 64->128, 3x3 512->512 with statistics + residual."""
    torch.manual_seed(0)
    dev = 'cuda'
    x = nhwc(torch.randn(8, 256, 64, 128, device=dev))
    w1 = ops.pack_conv_weight(torch.randn(512, 256, 1, 1, device=dev) * 0.05)
    bias = torch.randn(512, device=dev)
    x2 = nhwc(torch.randn(8, 64, 128, 256, device=dev))
    w2 = ops.pack_conv_weight(torch.randn(128, 64, 5, 5, device=dev) * 0.05)
    b2 = torch.randn(128, device=dev)
    x3 = nhwc(torch.randn(8, 512, 128, 256, device=dev))
    w3 = ops.pack_conv_weight(torch.randn(512, 512, 3, 3, device=dev) * 0.01)
    st = torch.zeros(2, 1, 512, device=dev)
    for _ in range(2):
        ops.conv2d_fprop(x, w1, 512, 1, 1, 1, 0, bias=bias, flags=ops.EPI_LRELU)
        ops.conv2d_fprop(x, w1, 512, 1, 1, 1, 0)
        ops.conv2d_fprop(x2, w2, 128, 5, 5, 1, 2, bias=b2, flags=ops.EPI_RELU)
        ops.conv2d_fprop(x3, w3, 512, 3, 3, 1, 1, stats=st, residual=x3)
    torch.cuda.synchronize()


if __name__ == '__main__':
    {'conv': conv, 'hbm': hbm, 'epi': epi, 'warp': warp}[sys.argv[1]]()
    print('ok')
