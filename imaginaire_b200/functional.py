"""Differentiable building blocks of the SPADE hot path on top of ops.py.

Each class is a ``torch.autograd.Function`` whose forward AND backward run the
hand-written kernels of libimx_kernels.so; PyTorch only keeps the graph.
Activations are bf16 NHWC (channels_last) tensors with the channel count padded
to a multiple of 8 (padding channels are zero).
"""
import math
import os

import torch
import torch.distributed as dist

from . import grad_sync, ops, peer_exchange

ACT_CODES = ops.ACT_CODES


# ------------------------------------------------------------ layout glue

class _ToNHWC(torch.autograd.Function):
    """NCHW float -> NHWC bf16 (channels zero-padded to a multiple of 8)."""

    @staticmethod
    def forward(ctx, x):
        ctx.c = x.shape[1]
        ctx.in_dtype = x.dtype
        return ops.nchw_f32_to_nhwc_bf16(x)

    @staticmethod
    def backward(ctx, g):
        return ops.nhwc_bf16_to_nchw_f32(_as_nhwc(g), ctx.c).to(ctx.in_dtype)


class _ToNCHW(torch.autograd.Function):
    """NHWC bf16 -> contiguous NCHW fp32 (first c channels), optional tanh."""

    @staticmethod
    def forward(ctx, x, c, apply_tanh):
        y = ops.nhwc_bf16_to_nchw_f32(x, c, apply_tanh)
        ctx.cp = x.shape[1]
        ctx.apply_tanh = apply_tanh
        if apply_tanh:
            ctx.save_for_backward(y)
        return y

    @staticmethod
    def backward(ctx, g):
        if ctx.apply_tanh:
            (y,) = ctx.saved_tensors
            g = g * (1.0 - y * y)
        return ops.nchw_f32_to_nhwc_bf16(g.contiguous(), ctx.cp), None, None


def _as_nhwc(t):
    """Makes a gradient tensor NHWC bf16 (autograd may hand back other layouts)."""
    if ops.is_nhwc_bf16(t):
        return t
    return t.to(torch.bfloat16).contiguous(memory_format=torch.channels_last)


def to_nhwc(x):
    """Any 4-D CUDA float tensor -> the internal activation layout."""
    if ops.is_nhwc_bf16(x) and x.shape[1] % 8 == 0:
        return x
    if x.dtype == torch.bfloat16 and x.shape[1] % 8 == 0:
        return x.contiguous(memory_format=torch.channels_last)
    return _ToNHWC.apply(x)


def to_nchw(x, channels=None, apply_tanh=False):
    return _ToNCHW.apply(x, x.shape[1] if channels is None else channels, apply_tanh)


class _Upsample2x(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        return ops.upsample2x(x)

    @staticmethod
    def backward(ctx, g):
        return ops.upsample2x_bwd(_as_nhwc(g))


class _BilinearUp2xAdd(torch.autograd.Function):
    @staticmethod
    def forward(ctx, src, lateral):
        ctx.has_lat = lateral is not None
        return ops.bilinear_up2x_add(src, lateral)

    @staticmethod
    def backward(ctx, g):
        g = _as_nhwc(g)
        return (ops.bilinear_up2x_bwd(g) if ctx.needs_input_grad[0] else None,
                g if ctx.has_lat and ctx.needs_input_grad[1] else None)


def bilinear_up2x_add(src, lateral=None):
    """up2x_bilinear(src) + lateral on NHWC bf16 tensors (the FPSE discriminator's top-down step)."""
    return _BilinearUp2xAdd.apply(to_nhwc(src), None if lateral is None else to_nhwc(lateral))


def upsample2x(x):
    y = _Upsample2x.apply(x)
    st = getattr(x, '_imx_stats', None)
    if st is not None:
        # every element is replicated 4 times: sums and count scale by 4, mean/var do not change
        sums, per_instance, count = st
        y._imx_stats = (sums * 4.0, per_instance, count * 4)
    return y


# ------------------------------------------------------------ convolution

_PACK_CACHE = {}


def _packed_weight(weight, inner_p, transpose_flip, inv_sigma=None):
    """bf16 GEMM-layout copy of a filter.  Parameters that are convolved with directly (no
    weight norm: frozen VGG filters, un-normalised SPADE projections) keep their packed copy
    until the optimizer changes them (tensor version counter); spectrally normalised filters
    are new tensors every forward and are packed each time."""
    # _imx_inner: the logical (unpadded) length of the contraction axis, for whoever counts algorithmic FLOPs
    inner = weight.shape[0] if transpose_flip else weight.shape[1]
    if inv_sigma is not None or not isinstance(weight, torch.nn.Parameter):
        packed = ops.pack_conv_weight(weight, inv_sigma, cin_p=inner_p, transpose_flip=transpose_flip)
        packed._imx_inner = inner
        return packed
    key = (id(weight), inner_p, transpose_flip)
    hit = _PACK_CACHE.get(key)
    if hit is not None and hit[0] == weight._version and hit[1] == weight.data_ptr():
        return hit[2]
    packed = ops.pack_conv_weight(weight, None, cin_p=inner_p, transpose_flip=transpose_flip)
    packed._imx_inner = inner
    _PACK_CACHE[key] = (weight._version, weight.data_ptr(), packed)
    return packed


def spectral_sigma(weight_orig, u_buf, v_buf, training, n_power_iterations=1, eps=1e-12, out=None):
    """Spectral norm of a filter with torch.nn.utils.spectral_norm's semantics (the reference's
    weight norm, imaginaire/layers/weight_norm.py:85): in training mode run the power iteration
    in place on the module's u / v buffers (no autograd), then sigma = u^T W v.  Returns
    (1/sigma as a 1-element tensor, u copy, v copy); the copies are what autograd may keep,
    because the buffers change again on the next forward.  Four launches of
    csrc/spectral_norm.cu instead of the hook's ~16 ATen / cuBLAS calls: W v of the last
    iteration is reused for sigma (torch recomputes the identical product), so the filter is read
    twice, and W / sigma is never materialised: the division happens inside the repack kernel."""
    with torch.no_grad():
        return ops.spectral_norm_sigma(weight_orig.detach(), u_buf, v_buf,
                                       n_power_iterations if training else 0, eps, out=out)


class _Conv2d(torch.autograd.Function):
    """y = act(conv(x, w * inv_sigma) + bias) (+ residual); x, y NHWC bf16; w fp32
    [Cout, Cin, KH, KW].  With spectral norm (inv_sigma, u, v given) `w` is weight_orig and
    the backward returns the gradient for weight_orig, including the path through sigma."""

    @staticmethod
    def forward(ctx, x, weight, bias, residual, stride, pad, act, stats, inv_sigma, u, v):
        cout, cin, kh, kw = weight.shape
        cin_p = x.shape[1]
        flags = {ops.ACT_NONE: 0, ops.ACT_LRELU: ops.EPI_LRELU, ops.ACT_RELU: ops.EPI_RELU}[act]
        if _small_cin(cin_p, kh, kw):
            # few input channels: patches (im2col) + a 1x1 GEMM with K = taps * 8
            col = ops.im2col_c8(x, kh, kw, stride, pad)
            wpk = _packed_weight(weight, 8, False, inv_sigma).view(cout, 1, 1, kh * kw * 8)
            y = ops.conv2d_fprop(col, wpk, cout, 1, 1, 1, 0, bias=bias, residual=residual, flags=flags, stats=stats)
        elif _small_cout(cout, kh, kw, stride, pad) and act == ops.ACT_NONE and residual is None and stats is None:
            # few output channels: per-tap products as a 1x1 GEMM, then the shifted sum over taps
            y = _small_cout_conv(x, weight.detach().permute(2, 3, 0, 1), bias, kh, kw, pad, cout, inv_sigma)
        else:
            # spectrally normalised parameter: the bf16 filter is packed unscaled once per optimizer step (cached on
            # the parameter's version) and 1 / sigma, which every application's power iteration changes, multiplies the
            # accumulator in the epilogue
            cached = inv_sigma is not None and isinstance(weight, torch.nn.Parameter)
            wpk = _packed_weight(weight, cin_p, False, None if cached else inv_sigma)
            y = ops.conv2d_fprop(x, wpk, cout, kh, kw, stride, pad, bias=bias, residual=residual, flags=flags,
                                 stats=stats, scale=inv_sigma if cached else None)
        ctx.stride, ctx.pad, ctx.act = stride, pad, act
        ctx.geom = (cout, cin, kh, kw, x.shape[2], x.shape[3])
        ctx.has_bias = bias is not None
        ctx.has_res = residual is not None
        needs_y = act != ops.ACT_NONE
        ctx.save_for_backward(x, weight, y if needs_y else None, inv_sigma, u, v)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, weight, y, inv_sigma, u, v = ctx.saved_tensors
        cout, cin, kh, kw, h, w = ctx.geom
        dy = _as_nhwc(dy)
        want_db = ctx.has_bias and ctx.needs_input_grad[2]
        chsum = getattr(dy, '_imx_chsum', None)     # left by the kernel that produced dy, if it had it for free
        if y is not None:
            dy, chsum = ops.act_bwd_bias(dy, y, ctx.act, want_db)
        dx = dw = db = dres = None
        if ctx.needs_input_grad[0]:
            dx = conv2d_dgrad(dy, weight, ctx.stride, ctx.pad, h, w, x.shape[1], inv_sigma)
        sync = grad_sync.lookup(weight) if (grad_sync._ACTIVE and ctx.needs_input_grad[1]) else None
        if sync is not None and not _small_cin(x.shape[1], kh, kw) and \
                not (dy.shape[1] == 8 and _small_cout(cout, kh, kw, ctx.stride, ctx.pad)):
            # data-parallel run: the split-K reduction also sums over the ranks (grad_sync.py); the
            # finished gradient is installed as weight.grad by PeerGradSync.finish()
            sync.wgrad(weight, x, dy, cin, cout, kh, kw, ctx.stride, ctx.pad,
                       sn=(weight.detach(), u, v, inv_sigma) if inv_sigma is not None else None)
        elif ctx.needs_input_grad[1]:
            # W_eff = W / sigma, sigma = u^T W v (u, v constants): dL/dW = G / sigma - (<G, W> / sigma^2) u v^T with
            # G = d/d(w * inv_sigma); the generic kernel forms <G, W> while it reduces its split-K partial sums
            sn = (weight.detach(), u, v, inv_sigma) if inv_sigma is not None else None
            dw = _conv2d_wgrad(x, dy, cin, cout, kh, kw, ctx.stride, ctx.pad, sn)
        if want_db:
            # over the padded channels, then the logical ones
            db = (chsum if chsum is not None else ops.channel_sum(dy))[:cout]
        if ctx.has_res and ctx.needs_input_grad[3]:
            dres = dy
        return dx, dw, db, dres, None, None, None, None, None, None, None


def _small_cin(cin_p, kh, kw):
    return cin_p == 8 and kh * kw > 1


def _small_cout(cout, kh, kw, stride, pad):
    return cout <= 8 and kh == kw and kh > 1 and stride == 1 and 2 * pad == kh - 1


def _small_cout_conv(x, w_taps, bias, kh, kw, pad, cout, inv_sigma):
    """'same' convolution with <= 8 output channels.  w_taps: (KH, KW, Cout, Cin) view of the filter.
    P[p][(tap, co)] = x[p] . w[tap, co, :] on the tensor cores (fp32 out), then the shifted tap sum."""
    cin = w_taps.shape[3]
    rows = w_taps.reshape(kh * kw * cout, cin, 1, 1)
    wpk = ops.pack_conv_weight(rows, inv_sigma, cin_p=x.shape[1])
    prod = ops.conv2d_fprop(x, wpk, kh * kw * cout, 1, 1, 1, 0, out_f32=True)
    return ops.shift_add_taps(prod, bias, kh, kw, pad, cout)


def _conv2d_wgrad(x, dy, cin, cout, kh, kw, stride, pad, sn=None):
    """Filter gradient [Cout, Cin, KH, KW] (fp32) of conv2d(x) given dy; with sn = (weight_orig, u, v, inv_sigma) the
    gradient for a spectrally normalised layer's weight_orig."""
    if _small_cin(x.shape[1], kh, kw):
        col = ops.im2col_c8(x, kh, kw, stride, pad)
        g = ops.conv2d_wgrad(col, dy, kh * kw * 8, cout, 1, 1, 1, 0)            # [Cout, (tap, ci)]
        dw = g.view(cout, kh, kw, 8).permute(0, 3, 1, 2)[:, :cin].contiguous()
    elif dy.shape[1] == 8 and _small_cout(cout, kh, kw, stride, pad):
        # dW[co][ci][t] = sum_p x[p][ci] * dy[p - t + pad][co]: patches of dy with flipped taps
        cin_p = x.shape[1]
        col = ops.im2col_c8(dy, kh, kw, 1, kh - 1 - pad)
        g = ops.conv2d_wgrad(col, x, kh * kw * 8, cin_p, 1, 1, 1, 0)            # [Cin_p, (flipped tap, co)]
        dw = g.view(cin_p, kh, kw, 8).flip(1, 2).permute(3, 0, 1, 2)[:cout, :cin].contiguous()
    else:
        return ops.conv2d_wgrad(x, dy, cin, cout, kh, kw, stride, pad, sn=sn)
    if sn is not None:
        ops.spectral_norm_weight_grad_(dw, *sn)
    return dw


def _parity_taps(k, pad, a):
    """Taps of a size-k, stride-2, padding-`pad` filter that reach input pixels of parity a (i = 2m + a):
    (first tap k0, number of taps T, left padding of the stride-1 correlation over dy), or None when no tap does or
    the padding would be negative.  dx[2m + a] = sum_t dy[m + t - left] * w[k0 + 2 (T - 1 - t)]."""
    k0 = (a + pad) % 2
    if k0 >= k:
        return None
    t = (k - k0 + 1) // 2
    left = (k0 + 2 * (t - 1) - a - pad) // 2
    return None if left < 0 else (k0, t, left)


def _dgrad_stride2(dy, weight, pad, h, w, cin_p, inv_sigma):
    """Data gradient of a stride-2 convolution as four stride-1 convolutions of dy, one per parity class of the input
    pixel, each with the filter taps of that parity, written straight into its pixels of dx
    (imx_conv2d_igemm_fprop_window_bf16).  A quarter of the multiply-adds of the zero-insertion form below, no
    zero-filled copy of dy.  The four transposed sub-filters are built once per optimizer step (two launches)."""
    cout, cin, kh, kw = weight.shape
    cout_p = dy.shape[1]
    plan = []
    for a in (0, 1):
        for b in (0, 1):
            ty, tx = _parity_taps(kh, pad, a), _parity_taps(kw, pad, b)
            if ty is None or tx is None:
                return None
            plan.append((a, b, ty, tx))
    is_param = isinstance(weight, torch.nn.Parameter)
    key = (id(weight), cout_p, 'dgrad2', pad)
    hit = _PACK_CACHE.get(key) if is_param else None
    if hit is not None and hit[0] == weight._version and hit[1] == weight.data_ptr():
        parts = hit[2]
    else:
        # one flip pack of the whole filter + one launch that sorts its taps into the four classes
        full = ops.pack_conv_weight(weight, None if is_param else inv_sigma, cin_p=cout_p, transpose_flip=True)
        parts = ops.split_dgrad_s2_filter(full, pad)
        for wpk in parts.values():
            wpk._imx_inner = cout
        if is_param:
            _PACK_CACHE[key] = (weight._version, weight.data_ptr(), parts)
    dx = ops.empty_nhwc(dy.shape[0], cin_p, h, w, dy.device)
    for a, b, (ky0, nty, py), (kx0, ntx, px) in plan:
        ma, mb = (h - a + 1) // 2, (w - b + 1) // 2
        if ma <= 0 or mb <= 0:
            continue
        ops.conv2d_fprop_window(dy, parts[(a, b)], cin, nty, ntx, py, px, ma, mb, dx, 2, 2, a, b,
                                scale=inv_sigma if is_param else None)
    return dx


def conv2d_dgrad(dy, weight, stride, pad, h, w, cin_p, inv_sigma=None):
    """Data gradient of conv2d as a stride-1 convolution of dy with the transposed,
    spatially flipped filter.  A stride-2 forward goes through _dgrad_stride2 (four parity-class convolutions);
    zero insertion remains for the few-channel layers and for geometries a parity class cannot express (a filter
    smaller than the stride, padding larger than the filter reaches)."""
    cout, cin, kh, kw = weight.shape
    cout_p = dy.shape[1]
    if kh != kw:
        raise RuntimeError('conv2d_dgrad: only square kernels are supported')
    if stride == 2 and not _small_cin(cout_p, kh, kw) and cin_p != 8 and os.environ.get('IMX_DGRAD_S2', '1') != '0':
        dx = _dgrad_stride2(dy, weight, pad, h, w, cin_p, inv_sigma)
        if dx is not None:
            return dx
    if stride != 1:
        n, _, ho, wo = dy.shape
        dh = (ho - 1) * stride + 1 + (h + 2 * pad - kh) % stride
        dw_ = (wo - 1) * stride + 1 + (w + 2 * pad - kw) % stride
        z = torch.zeros((n, dh, dw_, cout_p), dtype=torch.bfloat16, device=dy.device).permute(0, 3, 1, 2)
        z[:, :, 0:(ho - 1) * stride + 1:stride, 0:(wo - 1) * stride + 1:stride] = dy
        dy = z
    pad_t = kh - 1 - pad
    if _small_cin(cout_p, kh, kw):
        # dy has 8 channels (e.g. the to-RGB layer): patches of dy + a 1x1 GEMM
        col = ops.im2col_c8(dy, kh, kw, 1, pad_t)
        wt = _packed_weight(weight, 8, True, inv_sigma).view(cin, 1, 1, kh * kw * 8)
        return ops.conv2d_fprop(col, wt, cin, 1, 1, 1, 0, cout_p=cin_p)
    if cin_p == 8 and _small_cout(cin, kh, kw, 1, pad_t):
        # the input had 8 channels (image layers): per-tap products + shifted sum
        w_taps = weight.detach().flip(2, 3).permute(2, 3, 1, 0)    # (KH, KW, Cin, Cout) of the flipped filter
        return _small_cout_conv(dy, w_taps, None, kh, kw, pad_t, cin, inv_sigma)
    cached = inv_sigma is not None and isinstance(weight, torch.nn.Parameter)      # see _Conv2d.forward
    wt = _packed_weight(weight, cout_p, True, None if cached else inv_sigma)   # [Cin, KH, KW, Cout_p]
    return ops.conv2d_fprop(dy, wt, cin, kh, kw, 1, pad_t, cout_p=cin_p, scale=inv_sigma if cached else None)


def conv2d(x, weight, bias=None, residual=None, stride=1, pad=0, act=ops.ACT_NONE, stats_mode=None, sn=None):
    """sn = (inv_sigma, u, v) from spectral_sigma() when `weight` is a spectrally normalised
    layer's weight_orig.  stats_mode='batch' | 'instance': the epilogue also accumulates the per-channel sum / sum of squares
    of the stored output, attached to the result for the consuming norm layer."""
    stats = None
    if stats_mode in ('batch', 'instance'):
        g = x.shape[0] if stats_mode == 'instance' else 1
        stats = torch.zeros((2, g, weight.shape[0]), dtype=torch.float32, device=x.device)
    inv_sigma, u, v = sn if sn is not None else (None, None, None)
    y = _Conv2d.apply(x, weight, bias, residual, stride, pad, act, stats, inv_sigma, u, v)
    if stats is not None:
        n, _, ho, wo = y.shape
        y._imx_stats = (stats, stats_mode == 'instance', ho * wo if stats_mode == 'instance' else n * ho * wo)
    return y


# ---------------------------------------------------- normalisation + SPADE

class NormState:
    """What the fused kernels need from the inner norm module (parameters stay owned by
    the torch.nn norm module so state-dict keys are the reference's)."""

    def __init__(self, kind, weight, bias, running_mean, running_var, momentum, eps, training,
                 track_running_stats, num_batches_tracked=None, process_group=None):
        self.kind = kind                      # 'batch' | 'sync_batch' | 'instance'
        self.weight, self.bias = weight, bias
        self.running_mean, self.running_var = running_mean, running_var
        self.momentum, self.eps = momentum, eps
        self.training = training
        self.track_running_stats = track_running_stats
        self.num_batches_tracked = num_batches_tracked
        self.process_group = process_group

    @property
    def per_instance(self):
        return self.kind == 'instance'

    @property
    def use_batch_stats(self):
        return self.kind == 'instance' or self.training or self.running_mean is None


def _world(ns):
    if ns.kind != 'sync_batch' or not ns.training:
        return 1
    if not (dist.is_available() and dist.is_initialized()):
        return 1
    return dist.get_world_size(ns.process_group)


def compute_norm_stats(x, ns, up2x=False):
    """mean / rstd of x for the given inner norm (and the running-stat update).

    Statistics attached to x by the producing kernel's epilogue are used when present;
    otherwise the standalone statistics kernel runs.  SyncBatchNorm all-reduces
    [sum, sumsq, count] over the process group (NCCL on the compute stream).
    Returns (mean, rstd, count) with mean/rstd shaped (G*C,).
    """
    n, c, h, w = x.shape
    if not ns.use_batch_stats:
        rstd = torch.rsqrt(ns.running_var.float() + ns.eps)
        return ns.running_mean.float().contiguous(), rstd.contiguous(), n * h * w
    st = getattr(x, '_imx_stats', None)
    if st is not None and st[1] == ns.per_instance and st[0].shape[-1] == c:
        sums, _, count = st
    else:
        sums = ops.bn_stats(x, ns.per_instance)
        count = h * w if ns.per_instance else n * h * w
    world = _world(ns)
    if world > 1 and sums.is_cuda:
        # Equal per-rank batches (DistributedSampler; the backward assumes the same): the global count
        # is known on the host, so only [sum, sumsq] cross the wire and the single-GPU finalize kernel
        # does the rest (mean, rstd, running statistics) - 3 launches instead of ~15 small ATen ops.
        update = ns.track_running_stats and ns.running_mean is not None
        momentum = _momentum(ns) if update else 0.0
        ex = peer_exchange.get(ns.process_group) if sums.numel() <= peer_exchange.PeerExchange.MAX_FLOATS else None
        if ex is not None:
            # one single-CTA kernel over NVLink peer memory: push, flag, wait, sum in rank order, finalize
            mean, rstd = ex.bn_finalize(sums, count * world, ns.eps, ns.running_mean if update else None,
                                        ns.running_var if update else None, momentum)
            return mean, rstd, count * world
        gsums = sums.clone()          # the local sums may be shared with a sibling norm layer
        dist.all_reduce(gsums, op=dist.ReduceOp.SUM, group=ns.process_group)
        mean, rstd = ops.bn_finalize(gsums, count * world, ns.eps, ns.running_mean if update else None,
                                     ns.running_var if update else None, momentum)
        return mean, rstd, count * world
    if world > 1:
        mean, rstd, unbiased = sync_batch_stats(sums, count, ns.eps, ns.process_group)
        if ns.track_running_stats and ns.running_mean is not None:
            _update_running(ns, mean, unbiased)
        # the backward kernels take the global element count as a host scalar: equal
        # per-rank batches (DistributedSampler) are assumed there
        return mean, rstd, count * world
    update = ns.training and ns.track_running_stats and ns.running_mean is not None and not ns.per_instance
    momentum = _momentum(ns) if update else 0.0
    mean, rstd = ops.bn_finalize(sums, count, ns.eps, ns.running_mean if update else None,
                                 ns.running_var if update else None, momentum)
    return mean, rstd, count


def sync_batch_stats(sums, count, eps, group=None):
    """Cross-rank batch statistics from per-rank partial sums (works on any device / backend).

    sums: (2, 1, C) = [sum x, sum x^2] of this rank; count: this rank's elements per channel.
    One all-reduce of 2C+1 floats [sum, sumsq, count] replaces torch SyncBatchNorm's all-gather
    of [mean, invstd, count] + batch_norm_gather_stats_with_counts (torch
    nn/modules/_functions.py:49-86): same global mean / biased variance, also for unequal
    per-rank counts.  Returns (mean, rstd, unbiased_var), each (C,); nothing leaves the device."""
    # the count is written by a fill kernel: a Python scalar -> tensor copy would stall the launch queue
    packed = torch.cat([sums.reshape(-1), torch.full((1,), float(count), dtype=sums.dtype, device=sums.device)])
    dist.all_reduce(packed, op=dist.ReduceOp.SUM, group=group)
    gsums = packed[:-1].view_as(sums)
    count_t = packed[-1]
    mean = (gsums[0] / count_t).reshape(-1)
    var = (gsums[1].reshape(-1) / count_t - mean * mean).clamp_min_(0.0)
    rstd = torch.rsqrt(var + eps)
    return mean.contiguous(), rstd.contiguous(), var * (count_t / (count_t - 1.0))


def _momentum(ns):
    if ns.num_batches_tracked is not None:
        ns.num_batches_tracked.add_(1)
    if ns.momentum is None:   # cumulative moving average (calibrate_batch_norm_momentum)
        return 1.0 / float(ns.num_batches_tracked.item())
    return ns.momentum


def _update_running(ns, mean, unbiased_var):
    m = _momentum(ns)
    ns.running_mean.mul_(1.0 - m).add_(mean, alpha=m)
    ns.running_var.mul_(1.0 - m).add_(unbiased_var, alpha=m)


class _SpadeNormAct(torch.autograd.Function):
    """y = act(norm(x) * (1 + gamma) + beta) with materialised gamma/beta (NHWC bf16),
    per-sample gamma/beta vectors (AdaptiveNorm) or none (plain norm)."""

    @staticmethod
    def forward(ctx, x, weight, bias, gamma, beta, mean, rstd, count, per_instance, act, use_batch_stats, sync):
        y = ops.spade_modulate_fwd(x, mean, rstd, weight, bias, gamma, beta, per_instance, act)
        ctx.save_for_backward(x, weight, bias, gamma, beta, mean, rstd)
        ctx.cfg = (count, per_instance, act, use_batch_stats, sync)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, weight, bias, gamma, beta, mean, rstd = ctx.saved_tensors
        count, per_instance, act, use_batch_stats, sync = ctx.cfg
        hook = None
        if sync is not None:
            group, world = sync

            def hook(red):
                # SyncBatchNorm backward: all-reduce [sum dy', sum dy'*xn] (2C floats)
                ex = peer_exchange.get(group) if red.numel() <= peer_exchange.PeerExchange.MAX_FLOATS else None
                if ex is not None:
                    ex.allreduce_sum_(red)
                else:
                    dist.all_reduce(red, op=dist.ReduceOp.SUM, group=group)
                return count
        dx, dg, db, dwt, dbs = ops.spade_modulate_bwd(x, _as_nhwc(dy), mean, rstd, weight, bias, gamma, beta,
                                                      per_instance, act, count, use_batch_stats, hook)
        if not ctx.needs_input_grad[1]:
            dwt = None
        if not ctx.needs_input_grad[2]:
            dbs = None
        return dx, dwt, dbs, dg, db, None, None, None, None, None, None, None


def spade_norm_act(x, ns, gamma, beta, act):
    """Fused inner norm + modulation + activation (differentiable)."""
    mean, rstd, count = compute_norm_stats(x, ns)
    sync = None
    world = _world(ns)
    if world > 1:
        sync = (ns.process_group, world)
    w = ns.weight.float() if ns.weight is not None else None
    b = ns.bias.float() if ns.bias is not None else None
    return _SpadeNormAct.apply(x, w, b, gamma, beta, mean, rstd, count, ns.per_instance, act,
                               ns.use_batch_stats, sync)


def spade_gb_fused(hidden, w_gamma, w_beta, b_gamma, b_beta, x, ns, act, pad, inv_sigma_gamma=None,
                   inv_sigma_beta=None):
    """No-grad fast path: the gamma/beta projection conv applies the modulation in its
    epilogue, so gamma and beta never reach HBM."""
    mean, rstd, _ = compute_norm_stats(x, ns)
    c, ch, kh, kw = w_gamma.shape
    sg = sb = None
    if isinstance(w_gamma, torch.nn.Parameter) and isinstance(w_beta, torch.nn.Parameter) and \
            (inv_sigma_gamma is not None or inv_sigma_beta is not None):
        # spectrally normalised projections: the interleaved bf16 filter is packed unscaled once per optimizer step
        # (cached on the two parameters' versions); the two 1 / sigma multiply the accumulators in the epilogue
        key = (id(w_gamma), id(w_beta), hidden.shape[1], 'gb')
        hit = _PACK_CACHE.get(key)
        stamp = (w_gamma._version, w_beta._version, w_gamma.data_ptr(), w_beta.data_ptr())
        if hit is None or hit[0] != stamp:
            hit = (stamp, None, ops.pack_spade_gb_weight(w_gamma, w_beta, None, None, ch_p=hidden.shape[1]))
            _PACK_CACHE[key] = hit
        wgb = hit[2]
        sg = None if inv_sigma_gamma is None else inv_sigma_gamma.detach().float().reshape(1)
        sb = None if inv_sigma_beta is None else inv_sigma_beta.detach().float().reshape(1)
    else:
        wgb = ops.pack_spade_gb_weight(w_gamma, w_beta, inv_sigma_gamma, inv_sigma_beta, ch_p=hidden.shape[1])
    w = ns.weight.float() if ns.weight is not None else None
    b = ns.bias.float() if ns.bias is not None else None
    if w is None and b is not None:
        w = torch.ones_like(b)
    return ops.spade_gb_conv_modulate(hidden, wgb, b_gamma, b_beta, x, mean, rstd, w, b, c, kh, kw, pad,
                                      ns.per_instance, act, scale_gamma=sg, scale_beta=sb)


def _require_cuda(name, *ts):
    """The helpers below keep a torch expression for CUDA tensors whose dtype / layout the native
    kernels do not cover; there is no CPU path."""
    for t in ts:
        if not t.is_cuda:
            raise RuntimeError('%s: CUDA tensors required (this build has no CPU path)' % name)


# ------------------------------------------------------------------ feature-map L1 loss

class _L1Mean(torch.autograd.Function):
    @staticmethod
    def forward(ctx, a, b):
        ctx.save_for_backward(a, b)
        return (ops.l1_abs_sum(a, b) / a.numel()).reshape(())

    @staticmethod
    def backward(ctx, g):
        a, b = ctx.saved_tensors
        ga = ops.l1_abs_grad(a, b, g.detach().float().contiguous(), a.numel())
        return (ga if ctx.needs_input_grad[0] else None), (-ga if ctx.needs_input_grad[1] else None)


def l1_loss_mean(a, b):
    """``F.l1_loss(a.float(), b.float())`` (mean reduction) for feature maps.

    ref: nn.L1Loss in imaginaire/losses/perceptual.py:130-131 and feature_matching.py:36-38.
    bf16 CUDA tensors of one dense layout take two fused kernels (imx_l1_abs_sum_bf16 /
    imx_l1_abs_grad_bf16: 4 + 6 bytes per element instead of ~60 through ATen); the arithmetic is the
    same, |float(a) - float(b)| in fp32.  Anything else (fp32 inputs, mismatched layouts) goes
    through the torch expression."""
    _require_cuda('l1_loss_mean', a, b)
    if (a.is_cuda and b.is_cuda and a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16
            and ops._same_dense_layout(a, b) and a.numel() > 0):
        return _L1Mean.apply(a, b)
    return torch.nn.functional.l1_loss(a.float(), b.float())


# ------------------------------------------------------------------ 2x2 max pooling (VGG)

class _MaxPool2x2(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        ctx.save_for_backward(x)
        return ops.maxpool2x2(x)

    @staticmethod
    def backward(ctx, dy):
        x, = ctx.saved_tensors
        return ops.maxpool2x2_bwd(x, dy.contiguous(memory_format=torch.channels_last))


def max_pool2x2(x):
    """``F.max_pool2d(x, kernel_size=2, stride=2)``.  ref: the MaxPool2d(2, 2) stages of the perceptual
    loss' VGG-19 (imaginaire/losses/perceptual.py:175-192).  NHWC bf16 CUDA tensors with C % 8 == 0
    take the native kernels (no int64 index tensor; the backward recomputes the argmax); anything
    else goes through torch."""
    _require_cuda('max_pool2x2', x)
    if (x.is_cuda and x.dtype == torch.bfloat16 and x.dim() == 4 and x.shape[1] % 8 == 0
            and x.is_contiguous(memory_format=torch.channels_last) and x.numel() > 0):
        return _MaxPool2x2.apply(x)
    return torch.nn.functional.max_pool2d(x, kernel_size=2, stride=2)


# ------------------------------------------------------------------ 0.5x bilinear (discriminator input pyramid)

class _BilinearHalfAC(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, ho, wo):
        ctx.in_size = tuple(x.shape)
        ctx.out_size = (ho, wo)
        return ops.bilinear_resize_ac(x, ho, wo)

    @staticmethod
    def backward(ctx, g):
        # the (small) gradient goes through ATen's own backward kernel
        return torch.ops.aten.upsample_bilinear2d_backward(g.contiguous(), list(ctx.out_size), list(ctx.in_size), True,
                                                           None, None), None, None


def bilinear_half_align_corners(x):
    """``F.interpolate(x, scale_factor=0.5, mode='bilinear', align_corners=True)``.
    ref: imaginaire/discriminators/spade.py:86-88.  Contiguous NCHW fp32 CUDA tensors take the native
    forward kernel (ATen's arithmetic, one thread per output element); others go through torch."""
    _require_cuda('bilinear_half_align_corners', x)
    if x.is_cuda and x.dtype == torch.float32 and x.dim() == 4 and x.is_contiguous() and x.numel() > 0:
        ho, wo = int(math.floor(x.shape[2] * 0.5)), int(math.floor(x.shape[3] * 0.5))
        if ho > 0 and wo > 0:
            return _BilinearHalfAC.apply(x, ho, wo)
    return torch.nn.functional.interpolate(x, scale_factor=0.5, mode='bilinear', align_corners=True)


# ------------------------------------------------------------------ per-pixel channel dot (FPSE projection term)

class _ChannelDot(torch.autograd.Function):
    @staticmethod
    def forward(ctx, a, b):
        ctx.save_for_backward(a, b)
        return ops.channel_dot(a, b)

    @staticmethod
    def backward(ctx, g):
        a, b = ctx.saved_tensors
        return ops.channel_dot_bwd(a, b, g.to(torch.bfloat16).contiguous(), ctx.needs_input_grad[0], ctx.needs_input_grad[1])


def channel_dot(a, b):
    """``torch.mul(a, b).sum(dim=1, keepdim=True)``.  ref: imaginaire/discriminators/fpse.py:125-127.
    NHWC bf16 CUDA tensors of the same shape with C % 8 == 0 take one kernel each way (fp32
    accumulation, no product tensor); anything else goes through torch."""
    _require_cuda('channel_dot', a, b)

    def ok(t):
        return (t.is_cuda and t.dtype == torch.bfloat16 and t.dim() == 4 and t.shape[1] % 8 == 0
                and t.is_contiguous(memory_format=torch.channels_last))
    if a.shape == b.shape and ok(a) and ok(b) and a.numel() > 0:
        return _ChannelDot.apply(a, b)
    return torch.mul(a, b).sum(dim=1, keepdim=True)

