"""Tensor-level wrappers over the C ABI (include/imx.h).

PyTorch is plumbing here: it owns device memory and the current stream.  Every
function checks device / dtype / layout, allocates the outputs, and calls the
C entry point on ``torch.cuda.current_stream()``.  CPU tensors are rejected:
there is no CPU or eager fallback.

Activation layout: a logical N x C x H x W bf16 tensor in ``channels_last``
memory format, i.e. NHWC in memory, which is what the kernels read.
"""
import torch

from . import _C

EPI_BIAS, EPI_RELU, EPI_LRELU, EPI_RESIDUAL, EPI_TANH, EPI_STATS, EPI_OUT_F32 = 1, 2, 4, 8, 16, 32, 64
EPI_STATS_INSTANCE = 128
# layers with more output pixels than this always have enough tiles; skip the split-K query
_SPLITK_MAX_PIXELS = 148 * 256 * 2
ACT_NONE, ACT_LRELU, ACT_RELU = 0, 1, 2
ACT_CODES = {'': 0, 'none': 0, None: 0, 'leakyrelu': 1, 'relu': 2}


def _stream():
    # raw handle of torch's current stream; torch.cuda.current_stream() costs ~15 us of host time per call
    return torch._C._cuda_getCurrentRawStream(torch.cuda.current_device())


def _ptr(t):
    return None if t is None else t.data_ptr()


def _req_cuda(*ts):
    for t in ts:
        if t is not None and not t.is_cuda:
            raise RuntimeError('imaginaire_b200 ops run on CUDA tensors only (no CPU fallback); got a %s tensor'
                               % t.device.type)


def _req_f32c(t, name):
    if t.dtype != torch.float32 or not t.is_contiguous():
        raise RuntimeError('%s must be a contiguous float32 tensor' % name)


def round_up(v, m):
    return (v + m - 1) // m * m


def is_nhwc_bf16(t):
    return t.dim() == 4 and t.dtype == torch.bfloat16 and t.is_contiguous(memory_format=torch.channels_last)


def _req_nhwc(t, name):
    if not is_nhwc_bf16(t):
        raise RuntimeError('%s must be a 4-D bfloat16 tensor in channels_last (NHWC) memory; got %s %s strides %s'
                           % (name, t.dtype, tuple(t.shape), t.stride()))


def empty_nhwc(n, c, h, w, device, dtype=torch.bfloat16):
    """Logical (n, c, h, w) tensor with NHWC memory."""
    return torch.empty((n, h, w, c), dtype=dtype, device=device).permute(0, 3, 1, 2)


# --------------------------------------------------------------------- warp

def resample2d_fwd(img, flow, kernel_size=1):
    _req_cuda(img, flow)
    _req_f32c(img, 'input1'); _req_f32c(flow, 'input2')
    b, c, hi, wi = img.shape
    b2, two, h, w = flow.shape
    if two != 2 or b2 != b:
        raise RuntimeError('flow must be B x 2 x H x W with the same batch as the image')
    out = torch.empty((b, c, h, w), dtype=torch.float32, device=img.device)
    _C.lib().call('imx_resample2d_fwd_f32', _ptr(img), _ptr(flow), _ptr(out), b, c, h, w, hi, wi,
                  int(kernel_size), _stream())
    return out


def resample2d_bwd(img, flow, grad_out, kernel_size=1, trunc_alpha=True, need_img=True, need_flow=True):
    _req_cuda(img, flow, grad_out)
    _req_f32c(img, 'input1'); _req_f32c(flow, 'input2'); _req_f32c(grad_out, 'grad_output')
    b, c, hi, wi = img.shape
    _, _, h, w = flow.shape
    gimg = torch.zeros_like(img) if need_img else None
    gflow = torch.empty_like(flow) if need_flow else None
    _C.lib().call('imx_resample2d_bwd_f32', _ptr(img), _ptr(flow), _ptr(grad_out), _ptr(gimg), _ptr(gflow),
                  b, c, h, w, hi, wi, int(kernel_size), int(bool(trunc_alpha)), _stream())
    return gimg, gflow


def gridsample_warp_fwd(img, flow):
    _req_cuda(img, flow)
    _req_f32c(img, 'image'); _req_f32c(flow, 'flow')
    b, c, h, w = img.shape
    if tuple(flow.shape) != (b, 2, h, w):
        raise RuntimeError('flow must be B x 2 x H x W matching the image')
    out = torch.empty_like(img)
    _C.lib().call('imx_gridsample_warp_fwd_f32', _ptr(img), _ptr(flow), _ptr(out), b, c, h, w, _stream())
    return out


def gridsample_warp_bwd(img, flow, grad_out, need_img=True, need_flow=True):
    _req_cuda(img, flow, grad_out)
    _req_f32c(img, 'image'); _req_f32c(flow, 'flow'); _req_f32c(grad_out, 'grad_output')
    b, c, h, w = img.shape
    gimg = torch.zeros_like(img) if need_img else None
    gflow = torch.empty_like(flow) if need_flow else None
    _C.lib().call('imx_gridsample_warp_bwd_f32', _ptr(img), _ptr(flow), _ptr(grad_out), _ptr(gimg), _ptr(gflow),
                  b, c, h, w, _stream())
    return gimg, gflow


# ------------------------------------------------------------------- layout

def nearest_resize_to_nhwc(label, h, w, cp=None):
    """NCHW fp32 conditioning map -> nearest-resized NHWC bf16, channels padded to cp."""
    _req_cuda(label)
    label = label.float().contiguous()
    n, l, H, W = label.shape
    cp = round_up(l, 64) if cp is None else cp
    out = empty_nhwc(n, cp, h, w, label.device)
    _C.lib().call('imx_nearest_resize_nchw_f32_to_nhwc_bf16', _ptr(label), _ptr(out), n, l, H, W, h, w, cp, _stream())
    return out


def upsample2x(x):
    _req_cuda(x); _req_nhwc(x, 'x')
    n, c, h, w = x.shape
    out = empty_nhwc(n, c, 2 * h, 2 * w, x.device)
    _C.lib().call('imx_upsample2x_nhwc_bf16', _ptr(x), _ptr(out), n, h, w, c, _stream())
    return out


def upsample2x_bwd(g):
    _req_cuda(g); _req_nhwc(g, 'grad')
    n, c, h2, w2 = g.shape
    out = empty_nhwc(n, c, h2 // 2, w2 // 2, g.device)
    _C.lib().call('imx_upsample2x_bwd_nhwc_bf16', _ptr(g), _ptr(out), n, h2 // 2, w2 // 2, c, _stream())
    return out


def bilinear_up2x_add(src, lateral=None):
    """bilinear x2 upsampling (align_corners=False) of NHWC bf16 `src`, plus `lateral` if given."""
    _req_cuda(src); _req_nhwc(src, 'src')
    n, c, h, w = src.shape
    if lateral is not None:
        _req_nhwc(lateral, 'lateral')
        if tuple(lateral.shape) != (n, c, 2 * h, 2 * w):
            raise RuntimeError('lateral shape %s does not match the upsampled source %s' % (tuple(lateral.shape), (n, c, 2 * h, 2 * w)))
    out = empty_nhwc(n, c, 2 * h, 2 * w, src.device)
    _C.lib().call('imx_bilinear_up2x_add_nhwc_bf16', _ptr(src), _ptr(lateral), _ptr(out), n, h, w, c, _stream())
    return out


def bilinear_up2x_bwd(grad_out):
    _req_cuda(grad_out); _req_nhwc(grad_out, 'grad_out')
    n, c, hh, ww = grad_out.shape
    g = empty_nhwc(n, c, hh // 2, ww // 2, grad_out.device)
    _C.lib().call('imx_bilinear_up2x_bwd_nhwc_bf16', _ptr(grad_out), _ptr(g), n, hh // 2, ww // 2, c, _stream())
    return g


def nchw_f32_to_nhwc_bf16(x, cp=None):
    _req_cuda(x)
    x = x.float().contiguous()
    n, c, h, w = x.shape
    cp = round_up(c, 8) if cp is None else cp
    out = empty_nhwc(n, cp, h, w, x.device)
    _C.lib().call('imx_nchw_f32_to_nhwc_bf16', _ptr(x), _ptr(out), n, c, h, w, cp, _stream())
    return out


def nhwc_bf16_to_nchw_f32(x, c=None, apply_tanh=False):
    _req_cuda(x); _req_nhwc(x, 'x')
    n, cp, h, w = x.shape
    c = cp if c is None else c
    out = torch.empty((n, c, h, w), dtype=torch.float32, device=x.device)
    _C.lib().call('imx_nhwc_bf16_to_nchw_f32', _ptr(x), _ptr(out), n, c, h, w, cp, int(apply_tanh), _stream())
    return out


def pack_conv_weight(w, inv_sigma=None, cin_p=None, transpose_flip=False):
    """fp32 [Cout, Cin, KH, KW] -> bf16 [Cout, KH, KW, Cin_p] (or the dgrad filter
    [Cin, KH, KW, Cout_p] with transpose_flip)."""
    _req_cuda(w)
    w = w.detach().float().contiguous()
    cout, cin, kh, kw = w.shape
    inner = cout if transpose_flip else cin
    cin_p = round_up(inner, 8) if cin_p is None else cin_p
    rows = cin if transpose_flip else cout
    out = torch.empty((rows, kh, kw, cin_p), dtype=torch.bfloat16, device=w.device)
    if inv_sigma is not None:
        inv_sigma = inv_sigma.detach().float().reshape(1).contiguous()
    _C.lib().call('imx_pack_conv_weight_bf16', _ptr(w), _ptr(inv_sigma), _ptr(out), cout, cin, kh, kw, cin_p,
                  int(transpose_flip), _stream())
    return out


def split_dgrad_s2_filter(full, pad):
    """full: the transposed, flipped filter [Cin, KH, KW, Cout_p] (pack_conv_weight(..., transpose_flip=True)) of a
    stride-2 convolution.  Returns {(a, b): sub-filter [Cin, T(a), T(b), Cout_p]} for the four parity classes of its
    data gradient -- views of one buffer, one launch (imx_split_dgrad_s2_filter_bf16)."""
    _req_cuda(full)
    cin, kh, kw, cp = full.shape
    if full.dtype != torch.bfloat16 or not full.is_contiguous():
        raise RuntimeError('split_dgrad_s2_filter: full must be contiguous bf16 [Cin, KH, KW, Cout_p]')
    out = torch.empty(full.numel(), dtype=torch.bfloat16, device=full.device)
    _C.lib().call('imx_split_dgrad_s2_filter_bf16', _ptr(full), _ptr(out), cin, kh, kw, cp, pad, None, _stream())
    parts, off = {}, 0
    for a in (0, 1):
        ty = (kh - (a + pad) % 2 + 1) // 2
        for b in (0, 1):
            tx = (kw - (b + pad) % 2 + 1) // 2
            n = cin * ty * tx * cp
            parts[(a, b)] = out[off:off + n].view(cin, ty, tx, cp)
            off += n
    return parts


def pack_spade_gb_weight(w_gamma, w_beta, inv_sigma_gamma=None, inv_sigma_beta=None, ch_p=None):
    _req_cuda(w_gamma, w_beta)
    w_gamma = w_gamma.detach().float().contiguous()
    w_beta = w_beta.detach().float().contiguous()
    c, ch, kh, kw = w_gamma.shape
    ch_p = round_up(ch, 8) if ch_p is None else ch_p
    rows = 2 * round_up(c, 64)
    out = torch.empty((rows, kh, kw, ch_p), dtype=torch.bfloat16, device=w_gamma.device)
    isg = None if inv_sigma_gamma is None else inv_sigma_gamma.detach().float().reshape(1).contiguous()
    isb = None if inv_sigma_beta is None else inv_sigma_beta.detach().float().reshape(1).contiguous()
    _C.lib().call('imx_pack_spade_gb_weight_bf16', _ptr(w_gamma), _ptr(w_beta), _ptr(isg), _ptr(isb), _ptr(out),
                  c, ch, kh, kw, ch_p, _stream())
    return out


def spectral_norm_sigma(weight, u, v, n_power_iterations, eps, out=None):
    """In-place power iteration on the buffers u, v and sigma of `weight` viewed as
    [Cout, -1].  Returns (inv_sigma (1,), u copy, v copy)."""
    _req_cuda(weight, u, v)
    _req_f32c(weight, 'weight'); _req_f32c(u, 'u'); _req_f32c(v, 'v')
    r = weight.shape[0]
    c = weight.numel() // r
    lib = _C.lib()
    nbytes = lib.raw('imx_spectral_norm_scratch_bytes')(r, c)
    scratch = torch.empty((int(nbytes),), dtype=torch.uint8, device=weight.device)
    if out is None:
        out = torch.empty((1 + r + c,), dtype=torch.float32, device=weight.device)
    inv_sigma, u_copy, v_copy = out[:1], out[1:1 + r], out[1 + r:]
    lib.call('imx_spectral_norm_sigma_f32', _ptr(weight), _ptr(u), _ptr(v), _ptr(u_copy), _ptr(v_copy),
             _ptr(inv_sigma), _ptr(scratch), int(nbytes), r, c, int(n_power_iterations), float(eps), _stream())
    return inv_sigma, u_copy, v_copy


def spectral_norm_sigma_batched(descs, n_layers, max_r, max_c, scratch, out_floats, n_power_iterations, eps):
    """descs: int64 [n_layers, 8] device tensor (see include/imx.h).  Returns the output buffer."""
    out = torch.empty((out_floats,), dtype=torch.float32, device=descs.device)
    _C.lib().call('imx_spectral_norm_sigma_batched_f32', _ptr(descs), int(n_layers), int(max_r), int(max_c),
                  _ptr(scratch), _ptr(out), int(n_power_iterations), float(eps), _stream())
    return out


def spectral_norm_ema_update(descs, targets, n_layers, max_elems, sigma_out, beta):
    """targets[i] = beta * targets[i] + (1 - beta) * W_i / sigma_i for the layers of a batched sigma call."""
    _C.lib().call('imx_spectral_norm_ema_update_f32', _ptr(descs), _ptr(targets), int(n_layers), int(max_elems),
                  _ptr(sigma_out), float(beta), _stream())


def spectral_norm_weight_grad_(grad, weight, u, v, inv_sigma):
    """grad (= dL/d(W/sigma), fp32, same shape as weight) -> dL/dW, in place."""
    _req_cuda(grad, weight)
    _req_f32c(grad, 'grad'); _req_f32c(weight, 'weight')
    r = weight.shape[0]
    c = weight.numel() // r
    dot = torch.empty((1,), dtype=torch.float32, device=grad.device)
    _C.lib().call('imx_spectral_norm_weight_grad_f32', _ptr(grad), _ptr(weight), _ptr(u), _ptr(v), _ptr(inv_sigma),
                  _ptr(dot), r, c, _stream())
    return grad


# ------------------------------------------------------------- statistics

def bn_stats(x, per_instance=False):
    """Per-channel (or per-sample-per-channel) sum and sum of squares, fp32."""
    _req_cuda(x); _req_nhwc(x, 'x')
    n, c, h, w = x.shape
    g = n if per_instance else 1
    buf = torch.zeros((2, g, c), dtype=torch.float32, device=x.device)
    _C.lib().call('imx_bn_stats_nhwc_bf16', _ptr(x), _ptr(buf[0]), _ptr(buf[1]), n, h * w, c, int(per_instance),
                  _stream())
    return buf


def bn_finalize(sums, count, eps, running_mean=None, running_var=None, momentum=0.1):
    """sums: (2, G, C) from bn_stats (possibly all-reduced).  Returns (mean, rstd) as (G*C,)."""
    _req_cuda(sums)
    _, g, c = sums.shape
    out = torch.empty((2, g * c), dtype=torch.float32, device=sums.device)
    _C.lib().call('imx_bn_finalize', _ptr(sums[0]), _ptr(sums[1]), float(count), float(eps), _ptr(out[0]),
                  _ptr(out[1]), _ptr(running_mean), _ptr(running_var), float(momentum), g, c, _stream())
    return out[0], out[1]


def channel_sum(x, c=None):
    _req_cuda(x); _req_nhwc(x, 'x')
    n, cp, h, w = x.shape
    c = cp if c is None else c
    out = torch.zeros((c,), dtype=torch.float32, device=x.device)
    _C.lib().call('imx_channel_sum_nhwc_bf16', _ptr(x), _ptr(out), n * h * w, c, cp, _stream())
    return out


# -------------------------------------------------------------- modulation

def act_bwd_bias(dy, y, act, want_bias_grad):
    """(dy * act'(y), per-channel sum of it or None) for an activation fused into a conv epilogue."""
    _req_cuda(dy, y); _req_nhwc(dy, 'dy'); _req_nhwc(y, 'y')
    n, c, h, w = dy.shape
    dx = empty_nhwc(n, c, h, w, dy.device)
    bg = torch.zeros((c,), dtype=torch.float32, device=dy.device) if want_bias_grad else None
    _C.lib().call('imx_act_bwd_bias_nhwc_bf16', _ptr(dy), _ptr(y), _ptr(dx), _ptr(bg), n * h * w, c, int(act), _stream())
    return dx, bg


def bilinear_resize_ac(x, ho, wo):
    """F.interpolate(x, size=(ho, wo), mode='bilinear', align_corners=True) for a contiguous NCHW fp32 tensor."""
    _req_cuda(x)
    if x.dtype != torch.float32 or x.dim() != 4 or not x.is_contiguous():
        raise ValueError('bilinear_resize_ac: contiguous NCHW fp32 tensor expected')
    n, c, h, w = x.shape
    y = torch.empty((n, c, ho, wo), dtype=torch.float32, device=x.device)
    _C.lib().call('imx_bilinear_resize_ac_nchw_f32', _ptr(x), _ptr(y), n * c, h, w, ho, wo, _stream())
    return y


def channel_dot(a, b):
    """(N, 1, H, W) bf16: sum over the channels of a * b for two NHWC bf16 tensors."""
    _req_cuda(a, b); _req_nhwc(a, 'a'); _req_nhwc(b, 'b')
    n, c, h, w = a.shape
    out = torch.empty((n, 1, h, w), dtype=torch.bfloat16, device=a.device)
    _C.lib().call('imx_channel_dot_nhwc_bf16', _ptr(a), _ptr(b), _ptr(out), n * h * w, c, _stream())
    return out


def channel_dot_bwd(a, b, g, want_a=True, want_b=True):
    """Gradients of channel_dot: (g * b, g * a) as NHWC bf16 (None where not wanted); g: (N, 1, H, W) bf16 contiguous."""
    _req_cuda(a, b, g)
    n, c, h, w = a.shape
    da = empty_nhwc(n, c, h, w, a.device) if want_a else None
    db = empty_nhwc(n, c, h, w, a.device) if want_b else None
    _C.lib().call('imx_channel_dot_bwd_nhwc_bf16', _ptr(a), _ptr(b), _ptr(g), _ptr(da), _ptr(db), n * h * w, c, _stream())
    return da, db


def maxpool2x2(x):
    """2x2 / stride-2 max pooling of an NHWC bf16 tensor (floor mode)."""
    _req_cuda(x); _req_nhwc(x, 'x')
    n, c, h, w = x.shape
    y = empty_nhwc(n, c, h // 2, w // 2, x.device)
    _C.lib().call('imx_maxpool2x2_nhwc_bf16', _ptr(x), _ptr(y), n, h, w, c, _stream())
    return y


def maxpool2x2_bwd(x, dy):
    """Gradient of maxpool2x2 with respect to x (argmax recomputed from x)."""
    _req_cuda(x, dy); _req_nhwc(x, 'x'); _req_nhwc(dy, 'dy')
    n, c, h, w = x.shape
    dx = (torch.zeros if (h % 2 or w % 2) else torch.empty)((n, h, w, c), dtype=torch.bfloat16, device=x.device).permute(0, 3, 1, 2)
    _C.lib().call('imx_maxpool2x2_bwd_nhwc_bf16', _ptr(x), _ptr(dy), _ptr(dx), n, h, w, c, _stream())
    return dx


def _same_dense_layout(a, b):
    return (a.shape == b.shape and a.stride() == b.stride()
            and (a.is_contiguous() or (a.dim() == 4 and a.is_contiguous(memory_format=torch.channels_last))))


def l1_abs_sum(a, b):
    """1-element fp32 tensor: sum |a - b| over two bf16 tensors of one dense layout."""
    _req_cuda(a, b)
    if a.dtype != torch.bfloat16 or b.dtype != torch.bfloat16 or not _same_dense_layout(a, b):
        raise ValueError('l1_abs_sum: bf16 tensors of one dense memory layout expected')
    out = torch.zeros(1, dtype=torch.float32, device=a.device)
    _C.lib().call('imx_l1_abs_sum_bf16', _ptr(a), _ptr(b), a.numel(), _ptr(out), _stream())
    return out


def l1_abs_grad(a, b, upstream, denom):
    """bf16 tensor shaped and laid out like a: sign(a - b) * (upstream / denom) (upstream: fp32 device scalar)."""
    _req_cuda(a, b, upstream)
    if upstream.dtype != torch.float32 or upstream.numel() != 1:
        raise ValueError('l1_abs_grad: upstream must be one fp32 value on the device')
    grad = torch.empty_like(a)        # preserves the (dense) strides of a
    _C.lib().call('imx_l1_abs_grad_bf16', _ptr(a), _ptr(b), _ptr(grad), a.numel(), _ptr(upstream), float(denom), _stream())
    return grad


def _gb_mode(gamma):
    if gamma is None:
        return 2
    return 0 if gamma.dim() == 4 else 1


def spade_modulate_fwd(x, mean, rstd, w, b, gamma, beta, per_instance, act, up2x=False):
    _req_cuda(x); _req_nhwc(x, 'x')
    n, c, h, wd = x.shape
    if up2x:
        h, wd = 2 * h, 2 * wd
    mode = _gb_mode(gamma)
    if mode == 0:
        _req_nhwc(gamma, 'gamma'); _req_nhwc(beta, 'beta')
    elif mode == 1:
        _req_f32c(gamma, 'gamma'); _req_f32c(beta, 'beta')
    y = empty_nhwc(n, c, h, wd, x.device)
    _C.lib().call('imx_spade_modulate_fwd', _ptr(x), _ptr(mean), _ptr(rstd), _ptr(w), _ptr(b), _ptr(gamma),
                  _ptr(beta), _ptr(y), n, h, wd, c, int(per_instance), mode, int(act), int(up2x), _stream())
    return y


def spade_modulate_bwd(x, dy, mean, rstd, w, b, gamma, beta, per_instance, act, count, use_batch_stats=True,
                       reduce_hook=None):
    """Returns (dx, dgamma, dbeta, dweight, dbias).  ``reduce_hook(red)`` may all-reduce the
    (2, G, C) reductions in place across ranks (SyncBatchNorm backward) and return the
    global element count."""
    _req_cuda(x, dy); _req_nhwc(x, 'x'); _req_nhwc(dy, 'dy')
    n, c, h, wd = x.shape
    mode = _gb_mode(gamma)
    g = n if per_instance else 1
    red = torch.zeros((2, g, c), dtype=torch.float32, device=x.device)
    if mode == 0:
        dgamma = empty_nhwc(n, c, h, wd, x.device)
        dbeta = empty_nhwc(n, c, h, wd, x.device)
    elif mode == 1:
        dgamma = torch.zeros((n, c), dtype=torch.float32, device=x.device)
        dbeta = torch.zeros((n, c), dtype=torch.float32, device=x.device)
    else:
        dgamma = dbeta = None
    gb_sums = torch.zeros((2, c), dtype=torch.float32, device=x.device) if mode == 0 else None
    _C.lib().call('imx_spade_modulate_bwd_reduce', _ptr(x), _ptr(dy), _ptr(mean), _ptr(rstd), _ptr(w), _ptr(b),
                  _ptr(gamma), _ptr(beta), _ptr(dgamma), _ptr(dbeta), _ptr(red), _ptr(gb_sums), n, h, wd, c,
                  int(per_instance), mode, int(act), _stream())
    if gb_sums is not None:
        # picked up by the gamma / beta convolutions' backward as their bias gradients
        dgamma._imx_chsum = gb_sums[0]
        dbeta._imx_chsum = gb_sums[1]
    # dweight / dbias of the inner norm use the LOCAL reductions (DDP averages them later)
    dweight = red[1].sum(0) if w is not None else None
    dbias = red[0].sum(0) if b is not None else None
    if reduce_hook is not None and use_batch_stats:
        red = red.clone()
        count = reduce_hook(red)
    dx = empty_nhwc(n, c, h, wd, x.device)
    _C.lib().call('imx_spade_modulate_bwd_dx', _ptr(x), _ptr(dy), _ptr(mean), _ptr(rstd), _ptr(w), _ptr(b),
                  _ptr(gamma), _ptr(beta), _ptr(red), float(count), _ptr(dx), n, h, wd, c, int(per_instance),
                  mode, int(act), int(use_batch_stats), _stream())
    return dx, dgamma, dbeta, dweight, dbias


# ------------------------------------------------------------ convolutions

def conv2d_fprop_window(x, wpk, cout, kh, kw, pad_y, pad_x, ho, wo, out, stride_y, stride_x, y0, x0, scale=None):
    """Stride-1 convolution of x (NHWC bf16) whose ho x wo output pixels land at rows oy * stride_y + y0, columns
    ox * stride_x + x0 of the existing NHWC bf16 tensor `out` (imx_conv2d_igemm_fprop_window_bf16)."""
    _req_cuda(x, wpk, out); _req_nhwc(x, 'x'); _req_nhwc(out, 'out')
    n, cin_p, h, w = x.shape
    if wpk.dtype != torch.bfloat16 or not wpk.is_contiguous() or wpk.shape[-1] != cin_p:
        raise RuntimeError('packed weight must be contiguous bf16 [rows, KH, KW, Cin_p=%d]; got %s' %
                           (cin_p, tuple(wpk.shape)))
    if out.dtype != torch.bfloat16 or out.shape[0] != n:
        raise RuntimeError('out must be an NHWC bf16 tensor of the same batch size')
    if scale is not None:
        _req_f32c(scale, 'scale')
    _C.lib().call('imx_conv2d_igemm_fprop_window_bf16', _ptr(x), _ptr(wpk), _ptr(scale), None, None, _ptr(out),
                  n, h, w, cin_p, cout, out.shape[1], kh, kw, pad_y, pad_x, ho, wo, out.shape[2], out.shape[3],
                  stride_y, stride_x, y0, x0, 0, _stream())
    return out


def conv2d_fprop(x, wpk, cout, kh, kw, stride=1, pad=0, bias=None, residual=None, flags=0, stats=None,
                 cout_p=None, out_f32=False, scale=None):
    """x: NHWC bf16 (C = Cin_p), wpk: packed bf16 [Cout, KH, KW, Cin_p].  scale: optional 1-element fp32 device
    tensor; the accumulator is multiplied by it before the bias (1 / sigma with an unscaled packed filter)."""
    _req_cuda(x, wpk); _req_nhwc(x, 'x')
    n, cin_p, h, w = x.shape
    if wpk.dtype != torch.bfloat16 or not wpk.is_contiguous() or wpk.shape[-1] != cin_p:
        raise RuntimeError('packed weight must be contiguous bf16 [rows, KH, KW, Cin_p=%d]; got %s' %
                           (cin_p, tuple(wpk.shape)))
    cout_p = round_up(cout, 8) if cout_p is None else cout_p
    ho = (h + 2 * pad - kh) // stride + 1
    wo = (w + 2 * pad - kw) // stride + 1
    out = empty_nhwc(n, cout_p, ho, wo, x.device, torch.float32 if out_f32 else torch.bfloat16)
    if bias is not None:
        flags |= EPI_BIAS
        _req_f32c(bias, 'bias')
    if residual is not None:
        flags |= EPI_RESIDUAL
        _req_nhwc(residual, 'residual')
        if tuple(residual.shape) != (n, cout_p, ho, wo):
            raise RuntimeError('residual shape %s does not match the output %s' %
                               (tuple(residual.shape), (n, cout_p, ho, wo)))
    if stats is not None:
        # stats: zero-filled fp32 (2, G, Cout); G == N selects per-sample statistics
        flags |= EPI_STATS
        if stats.shape[1] != 1:
            if stats.shape[1] != n:
                raise RuntimeError('stats must be (2, 1, Cout) or (2, N, Cout)')
            flags |= EPI_STATS_INSTANCE
    if out_f32:
        flags |= EPI_OUT_F32
    lib = _C.lib()
    nbytes = 0
    if not (flags & EPI_STATS) and n * ho * wo <= _SPLITK_MAX_PIXELS:
        nbytes = lib.raw('imx_conv2d_fprop_workspace_bytes')(n, h, w, cin_p, cout, cout_p, kh, kw, stride, pad, flags)
        if nbytes < 0:
            raise _C.ImxError('imx_conv2d_fprop_workspace_bytes: %s' % lib.last_error())
    if scale is not None:
        _req_f32c(scale, 'scale')
        ws = torch.empty((int(nbytes),), dtype=torch.uint8, device=x.device) if nbytes > 0 else None
        lib.call('imx_conv2d_igemm_fprop_scaled_bf16', _ptr(x), _ptr(wpk), _ptr(scale), _ptr(bias), _ptr(residual), _ptr(out),
                 _ptr(stats), _ptr(ws), int(nbytes), n, h, w, cin_p, cout, cout_p, kh, kw, stride, pad, flags, _stream())
        if nbytes > 0:
            lib.launches += 1
    elif nbytes > 0:
        ws = torch.empty((int(nbytes),), dtype=torch.uint8, device=x.device)
        lib.call('imx_conv2d_igemm_fprop_splitk_bf16', _ptr(x), _ptr(wpk), _ptr(bias), _ptr(residual), _ptr(out),
                 _ptr(stats), _ptr(ws), int(nbytes), n, h, w, cin_p, cout, cout_p, kh, kw, stride, pad, flags, _stream())
    else:
        lib.call('imx_conv2d_igemm_fprop_bf16', _ptr(x), _ptr(wpk), _ptr(bias), _ptr(residual), _ptr(out),
                 _ptr(stats), n, h, w, cin_p, cout, cout_p, kh, kw, stride, pad, flags, _stream())
    return out


def spade_gb_conv_modulate(hidden, wgb, bias_gamma, bias_beta, x, mean, rstd, w, b, c, kh, kw, pad,
                           per_instance, act, up2x=False, scale_gamma=None, scale_beta=None):
    """scale_gamma / scale_beta: optional 1-element fp32 device tensors multiplying the gamma / beta accumulators
    (1 / sigma with an unscaled packed filter)."""
    _req_cuda(hidden, wgb, x); _req_nhwc(hidden, 'hidden'); _req_nhwc(x, 'x')
    n, ch_p, h, wd = hidden.shape
    y = empty_nhwc(n, c, h, wd, x.device)
    if scale_gamma is not None or scale_beta is not None:
        _C.lib().call('imx_spade_gb_conv_modulate_scaled_bf16', _ptr(hidden), _ptr(wgb), _ptr(scale_gamma), _ptr(scale_beta),
                      _ptr(bias_gamma), _ptr(bias_beta), _ptr(x), _ptr(mean), _ptr(rstd), _ptr(w), _ptr(b), _ptr(y), n, h, wd,
                      ch_p, c, kh, kw, pad, int(per_instance), int(act), int(up2x), _stream())
        return y
    _C.lib().call('imx_spade_gb_conv_modulate_bf16', _ptr(hidden), _ptr(wgb), _ptr(bias_gamma), _ptr(bias_beta),
                  _ptr(x), _ptr(mean), _ptr(rstd), _ptr(w), _ptr(b), _ptr(y), n, h, wd, ch_p, c, kh, kw, pad,
                  int(per_instance), int(act), int(up2x), _stream())
    return y


def im2col_c8(x, kh, kw, stride, pad):
    """x: NHWC bf16 with 8 channels -> (N, KH*KW*8, Ho, Wo) NHWC bf16 patches (tap-major, channel fastest)."""
    _req_cuda(x); _req_nhwc(x, 'x')
    n, c, h, w = x.shape
    if c != 8:
        raise RuntimeError('im2col_c8 needs an 8-channel tensor, got %d channels' % c)
    ho = (h + 2 * pad - kh) // stride + 1
    wo = (w + 2 * pad - kw) // stride + 1
    col = empty_nhwc(n, kh * kw * 8, ho, wo, x.device)
    _C.lib().call('imx_im2col_c8_nhwc_bf16', _ptr(x), _ptr(col), n, h, w, kh, kw, stride, pad, _stream())
    return col


def shift_add_taps(p, bias, kh, kw, pad, cout):
    """p: (N, ncols, H, W) fp32 NHWC per-tap products -> (N, 8, H, W) NHWC bf16 'same' convolution output."""
    _req_cuda(p)
    n, ncols, h, w = p.shape
    if p.dtype != torch.float32 or not p.is_contiguous(memory_format=torch.channels_last):
        raise RuntimeError('shift_add_taps: p must be fp32 in channels_last (NHWC) memory')
    if bias is not None:
        _req_f32c(bias, 'bias')
    y = empty_nhwc(n, 8, h, w, p.device)
    _C.lib().call('imx_shift_add_taps_f32', _ptr(p), _ptr(bias), _ptr(y), n, h, w, ncols, kh, kw, pad, cout, _stream())
    return y


def conv2d_wgrad(x, dy, cin, cout, kh, kw, stride=1, pad=0, out=None, sn=None):
    """dW (fp32, [Cout, Cin, KH, KW]) from x (NHWC bf16, C = Cin_p) and dy (NHWC bf16, C = Cout_p); written into `out`
    when given.  sn = (weight_orig, u, v, inv_sigma): the layer is spectrally normalised and the result is the
    gradient for weight_orig, G / sigma - (<G, W> / sigma^2) u v^T -- the inner product is formed while the split-K
    partial sums are reduced (imx_conv2d_igemm_wgrad_dot_bf16), the correction is one more pass."""
    _req_cuda(x, dy); _req_nhwc(x, 'x'); _req_nhwc(dy, 'dy')
    n, cin_p, h, w = x.shape
    cout_p = dy.shape[1]
    lib = _C.lib()
    nbytes = lib.raw('imx_conv2d_wgrad_workspace_bytes')(n, h, w, cin, cin_p, cout, cout_p, kh, kw, stride, pad)
    if nbytes < 0:
        raise _C.ImxError('imx_conv2d_wgrad_workspace_bytes: %s' % lib.last_error())
    ws = torch.empty((max(int(nbytes), 16),), dtype=torch.uint8, device=x.device)
    dw = out if out is not None else torch.empty((cout, cin, kh, kw), dtype=torch.float32, device=x.device)
    if sn is None or n == 0 or kh * kw > 47:
        lib.call('imx_conv2d_igemm_wgrad_bf16', _ptr(x), _ptr(dy), _ptr(dw), _ptr(ws), int(nbytes), n, h, w, cin, cin_p,
                 cout, cout_p, kh, kw, stride, pad, _stream())
        if sn is not None:
            spectral_norm_weight_grad_(dw, *sn)
        return dw
    weight, u, v, inv_sigma = sn
    _req_f32c(weight, 'weight')
    dot = torch.zeros((1,), dtype=torch.float32, device=x.device)
    lib.call('imx_conv2d_igemm_wgrad_dot_bf16', _ptr(x), _ptr(dy), _ptr(dw), _ptr(weight), _ptr(dot), _ptr(ws), int(nbytes),
             n, h, w, cin, cin_p, cout, cout_p, kh, kw, stride, pad, _stream())
    lib.call('imx_spectral_norm_weight_grad_apply_f32', _ptr(dw), _ptr(u), _ptr(v), _ptr(inv_sigma), _ptr(dot), cout,
             cin * kh * kw, _stream())
    return dw


def conv2d_wgrad_noalloc(x, dy, cin, cout, kh, kw, stride=1, pad=0, out=None):
    """conv2d_wgrad through the entry point that needs no workspace from the caller.

    For hosts that may not allocate device memory when the gradient is due (all memory
    reserved at start-up, allocation forbidden afterwards): the split-K scratch is a fixed
    array inside the library, so only the caller's own x, dy and the result buffer `out`
    (fp32 [Cout, Cin, KH, KW]; allocated here when not given) are involved.  Layers whose
    partial sums exceed the built-in array raise; calls must not overlap on the GPU."""
    _req_cuda(x, dy); _req_nhwc(x, 'x'); _req_nhwc(dy, 'dy')
    n, cin_p, h, w = x.shape
    cout_p = dy.shape[1]
    if out is None:
        out = torch.empty((cout, cin, kh, kw), dtype=torch.float32, device=x.device)
    _C.lib().call('imx_conv2d_igemm_wgrad_bf16_noalloc', _ptr(x), _ptr(dy), _ptr(out), n, h, w, cin, cin_p,
                  cout, cout_p, kh, kw, stride, pad, _stream())
    return out


# ------------------------------------------------------------ FlowNet2 operators

def _req_nchw_f32(t, name):
    if not t.is_cuda:
        raise RuntimeError('imaginaire_b200 ops run on CUDA tensors only (no CPU fallback); %s is on %s' % (name, t.device.type))
    if t.dim() != 4 or t.dtype != torch.float32 or not t.is_contiguous():
        raise RuntimeError('%s must be a contiguous 4-D float32 tensor' % name)


def channelnorm_fwd(x, norm_deg=2):
    _req_nchw_f32(x, 'input1')
    b, c, h, w = x.shape
    out = torch.empty((b, 1, h, w), dtype=torch.float32, device=x.device)
    _C.lib().call('imx_channelnorm_fwd_f32', _ptr(x), _ptr(out), b, c, h, w, int(norm_deg), _stream())
    return out


def channelnorm_bwd(x, out, grad_out, norm_deg=2):
    _req_nchw_f32(x, 'input1'); _req_nchw_f32(out, 'output'); _req_nchw_f32(grad_out, 'grad_output')
    b, c, h, w = x.shape
    gin = torch.empty_like(x)
    _C.lib().call('imx_channelnorm_bwd_f32', _ptr(x), _ptr(out), _ptr(grad_out), _ptr(gin), b, c, h, w, int(norm_deg), _stream())
    return gin


def correlation_output_size(c, h, w, pad_size, kernel_size, max_displacement, stride1, stride2):
    import ctypes
    oc, oh, ow = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
    _C.lib().call('imx_correlation_output_size', c, h, w, pad_size, kernel_size, max_displacement, stride1, stride2,
                  ctypes.byref(oc), ctypes.byref(oh), ctypes.byref(ow))
    _C.lib().launches -= 1      # host-side query, no kernel
    return oc.value, oh.value, ow.value


def _corr_scratch(x, pad_size):
    b, c, h, w = x.shape
    return torch.empty((b, h + 2 * pad_size, w + 2 * pad_size, c), dtype=torch.float32, device=x.device)


def correlation_fwd(x1, x2, pad_size, kernel_size, max_displacement, stride1, stride2):
    _req_nchw_f32(x1, 'input1'); _req_nchw_f32(x2, 'input2')
    if x1.shape != x2.shape:
        raise RuntimeError('correlation: the two inputs must have the same shape')
    b, c, h, w = x1.shape
    oc, oh, ow = correlation_output_size(c, h, w, pad_size, kernel_size, max_displacement, stride1, stride2)
    out = torch.empty((b, oc, oh, ow), dtype=torch.float32, device=x1.device)
    r1, r2 = _corr_scratch(x1, pad_size), _corr_scratch(x2, pad_size)
    lib = _C.lib()
    lib.call('imx_correlation_fwd_f32', _ptr(x1), _ptr(x2), _ptr(r1), _ptr(r2), _ptr(out), b, c, h, w, pad_size,
             kernel_size, max_displacement, stride1, stride2, _stream())
    lib.launches += 2
    return out


def correlation_bwd(x1, x2, grad_out, pad_size, kernel_size, max_displacement, stride1, stride2):
    _req_nchw_f32(x1, 'input1'); _req_nchw_f32(x2, 'input2'); _req_nchw_f32(grad_out, 'grad_output')
    b, c, h, w = x1.shape
    g1, g2 = torch.empty_like(x1), torch.empty_like(x2)
    r1, r2 = _corr_scratch(x1, pad_size), _corr_scratch(x2, pad_size)
    lib = _C.lib()
    lib.call('imx_correlation_bwd_f32', _ptr(x1), _ptr(x2), _ptr(grad_out), _ptr(r1), _ptr(r2), _ptr(g1), _ptr(g2),
             b, c, h, w, pad_size, kernel_size, max_displacement, stride1, stride2, _stream())
    lib.launches += 3
    return g1, g2
