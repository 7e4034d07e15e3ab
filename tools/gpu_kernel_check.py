"""Developer smoke check for the CUDA kernels against plain torch ops (NOT the parity
suite - that lives in tests/ and uses oracle/).  Each group runs in its own process so a
trapped kernel cannot poison the others:

    python tools/gpu_kernel_check.py            # all groups, one subprocess each
    python tools/gpu_kernel_check.py conv       # one group in-process
"""
import os
import subprocess
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

GROUPS = ['warp', 'layout', 'norm', 'conv_small', 'conv', 'conv_s2', 'gb', 'wgrad', 'modbwd', 'perf']


def rel(a, b):
    a = a.double(); b = b.double()
    return ((a - b).abs().max() / (b.abs().max() + 1e-30)).item()


def report(name, err, tol):
    print('%-58s err %.3e  tol %.1e  %s' % (name, err, tol, 'OK' if err <= tol else 'FAIL'), flush=True)
    return err <= tol


def bf16_round(t):
    return t.to('cuda').bfloat16().float()


def to_nhwc(t):
    """NCHW float -> logical NCHW bf16 with NHWC memory."""
    import torch
    return t.bfloat16().contiguous(memory_format=torch.channels_last)


def run_warp():
    import torch
    import torch.nn.functional as F
    from imaginaire_b200 import ops
    ok = True
    torch.manual_seed(0)
    for (B, C, H, W) in [(2, 3, 64, 128), (1, 3, 512, 1024), (2, 16, 33, 47)]:
        img = torch.randn(B, C, H, W, device='cuda')
        flow = F.interpolate(torch.randn(B, 2, max(H // 16, 2), max(W // 16, 2), device='cuda') * 8, size=(H, W),
                             mode='bilinear', align_corners=False).contiguous()
        out = ops.gridsample_warp_fwd(img, flow)
        # torch CPU reference of resample()
        ic, fc = img.cpu(), flow.cpu()
        gx = torch.linspace(-1, 1, W).view(1, 1, 1, W).expand(B, 1, H, W)
        gy = torch.linspace(-1, 1, H).view(1, 1, H, 1).expand(B, 1, H, W)
        grid = torch.cat([gx, gy], 1)
        fl = torch.cat([fc[:, 0:1] / ((W - 1.0) / 2.0), fc[:, 1:2] / ((H - 1.0) / 2.0)], 1)
        ref = F.grid_sample(ic, (grid + fl).permute(0, 2, 3, 1), mode='bilinear', padding_mode='border',
                            align_corners=True)
        d = (out.cpu() - ref).abs()
        print('  gridsample %s: maxabs %.3e, mismatching elements %d / %d' % ((B, C, H, W), d.max().item(),
                                                                             int((d > 0).sum()), d.numel()))
        ok &= report('gridsample_warp_fwd %s' % ((B, C, H, W),), rel(out.cpu(), ref), 1e-6)
        # resample2d forward vs a float64 torch restatement
        out2 = ops.resample2d_fwd(img, flow)
        xs = torch.arange(W, device='cuda').view(1, 1, W).float() + flow[:, 0]
        ys = torch.arange(H, device='cuda').view(1, H, 1).float() + flow[:, 1]
        fx, fy = torch.floor(xs), torch.floor(ys)
        a, b = (xs - fx).double(), (ys - fy).double()
        xL = fx.long().clamp(0, W - 1); xR = (fx + 1).long().clamp(0, W - 1)
        yT = fy.long().clamp(0, H - 1); yB = (fy + 1).long().clamp(0, H - 1)
        bi = torch.arange(B, device='cuda').view(B, 1, 1)
        ref2 = torch.zeros_like(img)
        for c in range(C):
            im = img[:, c].double()
            v = (1 - a) * (1 - b) * im[bi, yT, xL] + a * (1 - b) * im[bi, yT, xR] + \
                (1 - a) * b * im[bi, yB, xL] + a * b * im[bi, yB, xR]
            ref2[:, c] = v.float()
        ok &= report('resample2d_fwd %s' % ((B, C, H, W),), rel(out2, ref2), 1e-6)
        # backward vs autograd of grid_sample
        g = torch.randn_like(img)
        gi, gf = ops.gridsample_warp_bwd(img, flow, g)
        ic = ic.requires_grad_(True); fc = fc.requires_grad_(True)
        fl = torch.cat([fc[:, 0:1] / ((W - 1.0) / 2.0), fc[:, 1:2] / ((H - 1.0) / 2.0)], 1)
        r = F.grid_sample(ic, (grid + fl).permute(0, 2, 3, 1), mode='bilinear', padding_mode='border',
                          align_corners=True)
        r.backward(g.cpu())
        ok &= report('gridsample_warp_bwd d/dimg %s' % ((B, C, H, W),), rel(gi.cpu(), ic.grad), 1e-5)
        ok &= report('gridsample_warp_bwd d/dflow %s' % ((B, C, H, W),), rel(gf.cpu(), fc.grad), 1e-4)
        gi2, gf2 = ops.resample2d_bwd(img, flow, g, trunc_alpha=False)
        ok &= report('resample2d_bwd(floor) d/dimg vs gridsample', rel(gi2.cpu(), ic.grad), 2e-3)
        # informational only: the two warps clamp differently at the border (resample2d clamps tap indices,
        # grid_sample clips the coordinate and zeroes the gradient there), so large flows legitimately differ
        report('resample2d_bwd d/dflow vs gridsample (info)', rel(gf2.cpu(), fc.grad), float('inf'))
    return ok


def run_layout():
    import torch
    import torch.nn.functional as F
    from imaginaire_b200 import ops
    ok = True
    torch.manual_seed(0)
    lab = torch.randn(2, 35, 64, 96, device='cuda')
    for (h, w) in [(64, 96), (32, 48), (8, 12), (5, 7)]:
        out = ops.nearest_resize_to_nhwc(lab, h, w)
        ref = F.interpolate(lab, size=(h, w), mode='nearest').bfloat16().float()
        ok &= report('nearest_resize -> %dx%d' % (h, w), (out[:, :35].float() - ref).abs().max().item(), 0)
        ok &= report('  padding zero', out[:, 35:].float().abs().max().item(), 0)
    x = to_nhwc(torch.randn(2, 64, 6, 10, device='cuda'))
    up = ops.upsample2x(x)
    ok &= report('upsample2x', (up.float() - F.interpolate(x.float(), scale_factor=2, mode='nearest')).abs().max().item(), 0)
    g = to_nhwc(torch.randn(2, 64, 12, 20, device='cuda'))
    ub = ops.upsample2x_bwd(g)
    ok &= report('upsample2x_bwd', rel(ub.float(), F.avg_pool2d(g.float(), 2) * 4), 1e-2)
    t = torch.randn(3, 5, 7, 9, device='cuda')
    n = ops.nchw_f32_to_nhwc_bf16(t)
    ok &= report('nchw->nhwc', (n[:, :5].float() - t.bfloat16().float()).abs().max().item(), 0)
    back = ops.nhwc_bf16_to_nchw_f32(n, 5)
    ok &= report('nhwc->nchw', (back - t.bfloat16().float()).abs().max().item(), 0)
    w = torch.randn(24, 10, 3, 3, device='cuda')
    pk = ops.pack_conv_weight(w, torch.tensor([0.5], device='cuda'), cin_p=16)
    ref = torch.zeros(24, 3, 3, 16, device='cuda'); ref[..., :10] = (w * 0.5).permute(0, 2, 3, 1)
    ok &= report('pack_conv_weight', (pk.float() - ref.bfloat16().float()).abs().max().item(), 0)
    pt = ops.pack_conv_weight(w, None, cin_p=24, transpose_flip=True)
    ref = w.flip(2, 3).permute(1, 2, 3, 0)
    ok &= report('pack_conv_weight(transpose_flip)', (pt.float() - ref.bfloat16().float()).abs().max().item(), 0)
    return ok


def torch_spade(x, mean, rstd, w, b, gamma, beta, act):
    import torch.nn.functional as F
    xh = (x - mean) * rstd
    if w is not None:
        xh = xh * w + b
    y = xh * (1 + gamma) + beta if gamma is not None else xh
    if act == 1:
        y = F.leaky_relu(y, 0.2)
    elif act == 2:
        y = F.relu(y)
    return y


def run_norm():
    import torch
    from imaginaire_b200 import ops
    ok = True
    torch.manual_seed(0)
    for (N, C, H, W) in [(2, 64, 16, 32), (3, 512, 8, 8), (2, 2048, 4, 8), (1, 40, 5, 7), (2, 4096, 2, 2)]:
        x = to_nhwc(torch.randn(N, C, H, W, device='cuda') * 1.7 + 0.3)
        xf = x.float()
        for inst in (False, True):
            sums = ops.bn_stats(x, inst)
            dims = (2, 3) if inst else (0, 2, 3)
            ok &= report('bn_stats sum %s inst=%d' % ((N, C, H, W), inst), rel(sums[0].flatten(), xf.sum(dims).flatten()), 1e-5)
            ok &= report('bn_stats sumsq', rel(sums[1].flatten(), (xf * xf).sum(dims).flatten()), 1e-5)
            cnt = H * W if inst else N * H * W
            rm = torch.zeros(C, device='cuda'); rv = torch.ones(C, device='cuda')
            mean, rstd = ops.bn_finalize(sums, cnt, 1e-5, None if inst else rm, None if inst else rv, 0.1)
            m_ref = xf.mean(dims).flatten()
            v_ref = xf.var(dims, unbiased=False).flatten()
            ok &= report('bn_finalize mean', rel(mean, m_ref), 1e-5)
            ok &= report('bn_finalize rstd', rel(rstd, (v_ref + 1e-5).rsqrt()), 1e-4)
            if not inst:
                ok &= report('running_var', rel(rv, 0.9 + 0.1 * xf.var(dims, unbiased=True)), 1e-4)
            w = torch.randn(C, device='cuda'); b = torch.randn(C, device='cuda')
            gamma = to_nhwc(torch.randn(N, C, H, W, device='cuda') * 0.5)
            beta = to_nhwc(torch.randn(N, C, H, W, device='cuda') * 0.5)
            mv = mean.view(N, C, 1, 1) if inst else mean.view(1, C, 1, 1)
            rs = rstd.view(N, C, 1, 1) if inst else rstd.view(1, C, 1, 1)
            for act in (0, 1, 2):
                y = ops.spade_modulate_fwd(x, mean, rstd, w, b, gamma, beta, inst, act)
                ref = torch_spade(xf, mv, rs, w.view(1, C, 1, 1), b.view(1, C, 1, 1), gamma.float(), beta.float(), act)
                ok &= report('spade_modulate_fwd act=%d' % act, rel(y.float(), ref), 1e-2)
            gv = torch.randn(N, C, device='cuda'); bv = torch.randn(N, C, device='cuda')
            y = ops.spade_modulate_fwd(x, mean, rstd, None, None, gv, bv, inst, 1)
            ref = torch_spade(xf, mv, rs, None, None, gv.view(N, C, 1, 1), bv.view(N, C, 1, 1), 1)
            ok &= report('adaptive modulate', rel(y.float(), ref), 1e-2)
            y = ops.spade_modulate_fwd(x, mean, rstd, w, b, None, None, inst, 0)
            ref = torch_spade(xf, mv, rs, w.view(1, C, 1, 1), b.view(1, C, 1, 1), None, None, 0)
            ok &= report('plain norm', rel(y.float(), ref), 1e-2)
        if H % 2 == 0 and W % 2 == 0 and C % 8 == 0:
            xs = to_nhwc(torch.randn(N, C, H // 2, W // 2, device='cuda'))
            sums = ops.bn_stats(xs, False)
            mean, rstd = ops.bn_finalize(sums, N * H * W // 4, 1e-5)
            y = ops.spade_modulate_fwd(xs, mean, rstd, None, None, gamma, beta, False, 1, up2x=True)
            xu = torch.nn.functional.interpolate(xs.float(), scale_factor=2, mode='nearest')
            ref = torch_spade(xu, mean.view(1, C, 1, 1), rstd.view(1, C, 1, 1), None, None, gamma.float(), beta.float(), 1)
            ok &= report('modulate up2x', rel(y.float(), ref), 1e-2)
    return ok


def conv_case(N, Cin, Cout, H, W, k, stride, pad, flags=0, residual=False, stats=False, out_f32=False, cin_p=None):
    import torch
    import torch.nn.functional as F
    from imaginaire_b200 import ops
    cin_p = ops.round_up(Cin, 8) if cin_p is None else cin_p
    x = torch.randn(N, Cin, H, W, device='cuda')
    w = torch.randn(Cout, Cin, k, k, device='cuda') / (Cin * k * k) ** 0.5
    b = torch.randn(Cout, device='cuda')
    xp = torch.zeros(N, cin_p, H, W, device='cuda'); xp[:, :Cin] = x
    xn = to_nhwc(xp)
    wpk = ops.pack_conv_weight(w, None, cin_p=cin_p)
    cout_p = ops.round_up(Cout, 8)
    Ho = (H + 2 * pad - k) // stride + 1; Wo = (W + 2 * pad - k) // stride + 1
    res = to_nhwc(torch.randn(N, cout_p, Ho, Wo, device='cuda')) if residual else None
    st = torch.zeros(2, 1, Cout, device="cuda") if stats else None
    y = ops.conv2d_fprop(xn, wpk, Cout, k, k, stride, pad, bias=b, residual=res, flags=flags, stats=st, out_f32=out_f32)
    torch.cuda.synchronize()
    ref = F.conv2d(x.bfloat16().float(), w.bfloat16().float(), b, stride=stride, padding=pad)
    if residual:
        ref = ref + res[:, :Cout].float()
    if flags & ops.EPI_LRELU:
        ref = F.leaky_relu(ref, 0.2)
    if flags & ops.EPI_RELU:
        ref = F.relu(ref)
    if flags & ops.EPI_TANH:
        ref = torch.tanh(ref)
    name = 'conv N%d %d->%d %dx%d k%d s%d p%d f%d%s' % (N, Cin, Cout, H, W, k, stride, pad, flags, ' f32' if out_f32 else '')
    ok = report(name, rel(y[:, :Cout].float(), ref), 2e-3 if out_f32 else 1e-2)
    if cout_p > Cout:
        ok &= report('  padded channels zero', y[:, Cout:].float().abs().max().item(), 0)
    if stats:
        yq = y[:, :Cout].float()
        ok &= report('  epilogue stats sum', rel(st[0, 0], yq.sum((0, 2, 3))), 1e-3)
        ok &= report('  epilogue stats sumsq', rel(st[1, 0], (yq * yq).sum((0, 2, 3))), 1e-3)
    return ok


def run_conv_small():
    import torch
    torch.manual_seed(0)
    ok = conv_case(1, 64, 32, 8, 16, 1, 1, 0)          # one tile, one K step, BN=32
    ok &= conv_case(1, 64, 64, 8, 16, 3, 1, 1)         # 9 taps
    ok &= conv_case(1, 128, 128, 8, 16, 3, 1, 1)       # 2 K blocks per tap, BN=128
    ok &= conv_case(1, 64, 256, 8, 16, 1, 1, 0)        # BN=256
    return ok


def run_conv():
    import torch
    from imaginaire_b200 import ops
    torch.manual_seed(0)
    ok = True
    ok &= conv_case(2, 256, 512, 16, 32, 3, 1, 1)                       # 2 n-tiles, many m-tiles
    ok &= conv_case(2, 64, 64, 20, 24, 3, 1, 1)                         # ragged tiles (W not a power of two)
    ok &= conv_case(1, 40, 72, 13, 11, 5, 1, 2)                         # odd channels / dims, 5x5
    ok &= conv_case(2, 192, 64, 16, 16, 5, 1, 2, flags=ops.EPI_RELU)    # label->hidden shape
    ok &= conv_case(2, 128, 256, 32, 64, 3, 1, 1, flags=ops.EPI_LRELU, residual=True)
    ok &= conv_case(1, 256, 3, 64, 64, 5, 1, 2, flags=ops.EPI_TANH)     # to-RGB
    ok &= conv_case(2, 128, 128, 16, 32, 3, 1, 1, stats=True)
    ok &= conv_case(2, 128, 128, 16, 32, 3, 1, 1, out_f32=True)
    ok &= conv_case(4, 512, 1024, 32, 64, 3, 1, 1)                      # > 148 tiles: persistent loop + TMEM double buffering
    ok &= conv_case(1, 1024, 1024, 16, 32, 3, 1, 1)
    ok &= conv_case(1, 64, 64, 256, 512, 1, 1, 0)
    return ok


def run_conv_s2():
    import torch
    torch.manual_seed(0)
    ok = conv_case(2, 64, 64, 32, 32, 3, 2, 1)
    ok &= conv_case(1, 8, 32, 256, 256, 3, 2, 1)        # style-encoder layer1 shape (3 -> 8 padded)
    ok &= conv_case(2, 64, 128, 64, 64, 4, 2, 2)        # discriminator shape (k4 s2 p2)
    ok &= conv_case(2, 64, 128, 33, 47, 4, 2, 1)
    return ok


def run_gb():
    import torch
    import torch.nn.functional as F
    from imaginaire_b200 import ops
    torch.manual_seed(0)
    ok = True
    for (N, Ch, C, H, W, k, inst, up) in [(2, 64, 64, 16, 32, 5, False, False), (2, 128, 512, 16, 32, 5, False, False),
                                          (1, 64, 256, 16, 16, 3, True, False), (2, 128, 128, 32, 32, 5, False, True),
                                          (2, 64, 96, 12, 20, 3, True, False)]:
        hid = torch.relu(torch.randn(N, Ch, H, W, device='cuda'))
        wg = torch.randn(C, Ch, k, k, device='cuda') / (Ch * k * k) ** 0.5
        wb = torch.randn(C, Ch, k, k, device='cuda') / (Ch * k * k) ** 0.5
        bg = torch.randn(C, device='cuda') * 0.1; bb = torch.randn(C, device='cuda') * 0.1
        hx, wx = (H // 2, W // 2) if up else (H, W)
        x = to_nhwc(torch.randn(N, C, hx, wx, device='cuda') * 1.5 + 0.2)
        sums = ops.bn_stats(x, inst)
        mean, rstd = ops.bn_finalize(sums, hx * wx if inst else N * hx * wx, 1e-5)
        nw = torch.randn(C, device='cuda'); nb = torch.randn(C, device='cuda')
        wgb = ops.pack_spade_gb_weight(wg, wb)
        y = ops.spade_gb_conv_modulate(to_nhwc(hid), wgb, bg, bb, x, mean, rstd, nw, nb, C, k, k, k // 2, inst, 1, up)
        torch.cuda.synchronize()
        hq = hid.bfloat16().float()
        gamma = F.conv2d(hq, wg.bfloat16().float(), bg, padding=k // 2)
        beta = F.conv2d(hq, wb.bfloat16().float(), bb, padding=k // 2)
        xf = x.float()
        if up:
            xf = F.interpolate(xf, scale_factor=2, mode='nearest')
        mv = mean.view(N, C, 1, 1) if inst else mean.view(1, C, 1, 1)
        rs = rstd.view(N, C, 1, 1) if inst else rstd.view(1, C, 1, 1)
        ref = torch_spade(xf, mv, rs, nw.view(1, C, 1, 1), nb.view(1, C, 1, 1), gamma, beta, 1)
        ok &= report('spade_gb_conv_modulate %s' % ((N, Ch, C, H, W, k, inst, up),), rel(y.float(), ref), 1e-2)
    return ok


def run_wgrad():
    import torch
    import torch.nn.functional as F
    from imaginaire_b200 import ops
    torch.manual_seed(0)
    ok = True
    for (N, Cin, Cout, H, W, k, s, p) in [(1, 64, 128, 8, 8, 1, 1, 0), (2, 64, 64, 16, 32, 3, 1, 1),
                                          (2, 128, 256, 16, 32, 3, 1, 1), (2, 256, 128, 16, 16, 3, 1, 1),
                                          (1, 40, 72, 13, 11, 5, 1, 2), (2, 64, 512, 32, 64, 5, 1, 2),
                                          (2, 64, 128, 32, 32, 4, 2, 1), (4, 512, 512, 32, 64, 3, 1, 1)]:
        cin_p = ops.round_up(Cin, 8); cout_p = ops.round_up(Cout, 8)
        Ho = (H + 2 * p - k) // s + 1; Wo = (W + 2 * p - k) // s + 1
        x = torch.zeros(N, cin_p, H, W, device='cuda'); x[:, :Cin] = torch.randn(N, Cin, H, W, device='cuda')
        dy = torch.zeros(N, cout_p, Ho, Wo, device='cuda'); dy[:, :Cout] = torch.randn(N, Cout, Ho, Wo, device='cuda')
        dw = ops.conv2d_wgrad(to_nhwc(x), to_nhwc(dy), Cin, Cout, k, k, s, p)
        torch.cuda.synchronize()
        xq = x[:, :Cin].bfloat16().float(); dq = dy[:, :Cout].bfloat16().float()
        ref = torch.nn.grad.conv2d_weight(xq, (Cout, Cin, k, k), dq, stride=s, padding=p)
        ok &= report('wgrad %s' % ((N, Cin, Cout, H, W, k, s, p),), rel(dw, ref), 2e-3)
    return ok


def run_modbwd():
    import torch
    from imaginaire_b200 import ops
    torch.manual_seed(0)
    ok = True
    for (N, C, H, W, inst, mode, act, affine) in [(2, 64, 8, 16, False, 0, 1, True), (2, 64, 8, 16, True, 0, 1, True),
                                                  (3, 128, 4, 4, False, 1, 1, False), (2, 64, 8, 8, False, 2, 0, True),
                                                  (2, 512, 8, 8, True, 0, 0, False)]:
        x = to_nhwc(torch.randn(N, C, H, W, device='cuda') * 1.5 + 0.5)
        dy = to_nhwc(torch.randn(N, C, H, W, device='cuda'))
        w = (torch.randn(C, device='cuda') + 1.5) if affine else None
        b = torch.randn(C, device='cuda') if affine else None
        if mode == 0:
            gamma = to_nhwc(torch.randn(N, C, H, W, device='cuda') * 0.5); beta = to_nhwc(torch.randn(N, C, H, W, device='cuda'))
        elif mode == 1:
            gamma = torch.randn(N, C, device='cuda') * 0.5; beta = torch.randn(N, C, device='cuda')
        else:
            gamma = beta = None
        sums = ops.bn_stats(x, inst)
        cnt = H * W if inst else N * H * W
        mean, rstd = ops.bn_finalize(sums, cnt, 1e-5)
        dx, dg, db, dwt, dbs = ops.spade_modulate_bwd(x, dy, mean, rstd, w, b, gamma, beta, inst, act, cnt)
        # autograd reference in fp32 on the same (bf16-valued) inputs
        xr = x.float().requires_grad_(True)
        wr = w.clone().requires_grad_(True) if affine else None
        br = b.clone().requires_grad_(True) if affine else None
        dims = (2, 3) if inst else (0, 2, 3)
        m = xr.mean(dims, keepdim=True); v = xr.var(dims, unbiased=False, keepdim=True)
        xh = (xr - m) * (v + 1e-5).rsqrt()
        if affine:
            xh = xh * wr.view(1, C, 1, 1) + br.view(1, C, 1, 1)
        if mode == 0:
            gr = gamma.float().requires_grad_(True); ber = beta.float().requires_grad_(True)
            pre = xh * (1 + gr) + ber
        elif mode == 1:
            gr = gamma.clone().requires_grad_(True); ber = beta.clone().requires_grad_(True)
            pre = xh * (1 + gr.view(N, C, 1, 1)) + ber.view(N, C, 1, 1)
        else:
            pre = xh
        y = torch.nn.functional.leaky_relu(pre, 0.2) if act == 1 else pre
        y.backward(dy.float())
        tag = str((N, C, H, W, inst, mode, act, affine))
        ok &= report('modulate_bwd dx %s' % tag, rel(dx.float(), xr.grad), 2e-2)
        if mode != 2:
            ok &= report('  dgamma', rel(dg.float(), gr.grad), 1e-2)
            ok &= report('  dbeta', rel(db.float(), ber.grad), 1e-2)
        if affine:
            ok &= report('  dweight', rel(dwt, wr.grad), 1e-3)
            ok &= report('  dbias', rel(dbs, br.grad), 1e-3)
    return ok


def run_perf():
    import torch
    from imaginaire_b200 import ops

    def timeit(fn, iters=10):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        s = torch.cuda.Event(enable_timing=True); e = torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(iters):
            fn()
        e.record(); torch.cuda.synchronize()
        return s.elapsed_time(e) / iters

    torch.manual_seed(0)
    for (N, Cin, Cout, H, W, k) in [(8, 512, 512, 128, 256, 3), (8, 2048, 2048, 16, 32, 3), (8, 128, 512, 128, 256, 5),
                                    (8, 1024, 1024, 32, 64, 3), (8, 64, 128, 128, 256, 5)]:
        x = to_nhwc(torch.randn(N, Cin, H, W, device='cuda'))
        wpk = ops.pack_conv_weight(torch.randn(Cout, Cin, k, k, device='cuda') * 0.01)
        ms = timeit(lambda: ops.conv2d_fprop(x, wpk, Cout, k, k, 1, k // 2))
        fl = 2.0 * N * H * W * Cout * Cin * k * k
        print('conv fprop N%d %d->%d %dx%d k%d: %.3f ms  %.1f TFLOP/s' % (N, Cin, Cout, H, W, k, ms, fl / ms / 1e9), flush=True)
        xt = x.float().contiguous(memory_format=torch.channels_last).bfloat16()
        wt = torch.randn(Cout, Cin, k, k, device='cuda').bfloat16().contiguous(memory_format=torch.channels_last)
        ms2 = timeit(lambda: torch.nn.functional.conv2d(xt, wt, padding=k // 2))
        print('   cuDNN bf16 channels_last: %.3f ms  %.1f TFLOP/s' % (ms2, fl / ms2 / 1e9), flush=True)
        dy = to_nhwc(torch.randn(N, Cout, H, W, device='cuda'))
        ms3 = timeit(lambda: ops.conv2d_wgrad(x, dy, Cin, Cout, k, k, 1, k // 2))
        print('   wgrad: %.3f ms  %.1f TFLOP/s' % (ms3, fl / ms3 / 1e9), flush=True)
    for (N, C, H, W) in [(8, 512, 128, 256), (8, 2048, 16, 32), (8, 256, 256, 512)]:
        x = to_nhwc(torch.randn(N, C, H, W, device='cuda'))
        g = to_nhwc(torch.randn(N, C, H, W, device='cuda')); b = to_nhwc(torch.randn(N, C, H, W, device='cuda'))
        nbytes = x.numel() * 2
        ms = timeit(lambda: ops.bn_stats(x, False))
        print('bn_stats %s: %.3f ms  %.0f GB/s' % ((N, C, H, W), ms, nbytes / ms / 1e6), flush=True)
        sums = ops.bn_stats(x, False)
        mean, rstd = ops.bn_finalize(sums, N * H * W, 1e-5)
        ms = timeit(lambda: ops.spade_modulate_fwd(x, mean, rstd, None, None, g, b, False, 1))
        print('spade_modulate_fwd: %.3f ms  %.0f GB/s' % (ms, 4 * nbytes / ms / 1e6), flush=True)
    for (B, C, H, W) in [(1, 3, 512, 1024), (8, 3, 512, 1024), (8, 64, 512, 1024)]:
        img = torch.randn(B, C, H, W, device='cuda'); flow = torch.randn(B, 2, H, W, device='cuda') * 4
        ms = timeit(lambda: ops.resample2d_fwd(img, flow))
        print('resample2d_fwd %s: %.4f ms  %.0f GB/s' % ((B, C, H, W), ms, (2 * C + 2) * B * H * W * 4 / ms / 1e6), flush=True)
        ms = timeit(lambda: ops.gridsample_warp_fwd(img, flow))
        print('gridsample_warp_fwd: %.4f ms  %.0f GB/s' % (ms, (2 * C + 2) * B * H * W * 4 / ms / 1e6), flush=True)
    return True


if __name__ == '__main__':
    if len(sys.argv) > 1:
        ok = globals()['run_' + sys.argv[1]]()
        print('GROUP %s: %s' % (sys.argv[1], 'PASS' if ok else 'FAIL'))
        sys.exit(0 if ok else 1)
    results = {}
    for g in GROUPS:
        t0 = time.time()
        try:
            r = subprocess.run([sys.executable, os.path.abspath(__file__), g], timeout=300)
            results[g] = 'PASS' if r.returncode == 0 else 'FAIL(rc=%d)' % r.returncode
        except subprocess.TimeoutExpired:
            results[g] = 'TIMEOUT'
        print('=== %s: %s (%.1fs)' % (g, results[g], time.time() - t0), flush=True)
    print(results)
