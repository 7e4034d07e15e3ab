"""CPU oracle of the SPADE hot path -- TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

A functional, fp32, CPU restatement of the reference's SPADE algorithm, driven by a
reference-format ``state_dict`` (same keys as ``imaginaire.generators.spade.Generator``).
Every function cites the reference lines it follows.  The convolution / linear products
themselves are torch's CPU ``F.conv2d`` / ``F.linear`` (the reference's own arithmetic
lives in PyTorch, a third-party dependency: SURVEY.md section 8(c)); everything the
reference composes around them -- nearest resize indexing, spectral normalisation,
instance / batch statistics with running-stat updates, modulation, activations, block
and network structure -- is written out explicitly here.

Pinned against the reference itself: ``oracle/make_golden.py`` runs the reference's own
modules (imported from /root/reference in the build container) on seeded inputs and
stores their outputs under ``tests/golden/``; ``tests/test_oracle.py`` checks this file
against those fixtures.  The reference has no tests or golden vectors of its own
(SURVEY.md section 4): parity is unpinned BY the reference and pinned by these fixtures.
"""
import math

import torch
import torch.nn.functional as F

EPS_BN = 1e-5

# Storage-precision emulation.  None (default): the fp32 oracle.  Set to a function (e.g.
# ``lambda t: t.bfloat16().float()``) to round conv operands (activations and weights) and
# every stored activation the way a bf16-storage / fp32-accumulate pipeline does; the tests
# use this to measure how much of the GPU-vs-fp32 difference is inherent to bf16 storage.
QUANT = None


def _q(t):
    return t if QUANT is None else QUANT(t)


# Second emulation level: STORE (same kind of function) rounds every tensor the native pipeline keeps in HBM between
# two kernels -- convolution outputs (after bias / activation / residual sum), gamma and beta when autograd
# materialises them (STORE_GAMMA_BETA) -- on top of QUANT's rounding of the convolution operands.  QUANT alone is "one
# rounding per convolution input"; QUANT + STORE is the rounding schedule of a bf16-storage pipeline in which the
# normalisation statistics are not known when a convolution's output is written, so that output is rounded once when
# stored and once more after the modulation.
STORE = None
STORE_GAMMA_BETA = True

# Debugging aid: set to a list to receive (block name, output) after every block of spade_generator().
TRACE = None


def _s(t):
    return t if STORE is None else STORE(t)


def _trace(name, t):
    if TRACE is not None:
        TRACE.append((name, t.detach().clone()))
    return t


# --------------------------------------------------------------- primitives

def nearest_resize(x, size):
    """F.interpolate(x, size, mode='nearest'): src = floor(dst * in / out), scale in fp32.
    ref: imaginaire/layers/activation_norm.py:224-225, generators/spade.py:424"""
    h, w = size
    H, W = x.shape[2], x.shape[3]
    sy = torch.tensor(H / h, dtype=torch.float32)
    sx = torch.tensor(W / w, dtype=torch.float32)
    iy = torch.floor(torch.arange(h, dtype=torch.float32) * sy).long().clamp_(max=H - 1)
    ix = torch.floor(torch.arange(w, dtype=torch.float32) * sx).long().clamp_(max=W - 1)
    return x[:, :, iy][:, :, :, ix]


def upsample2x(x):
    """NearestUpsample(scale_factor=2). ref: imaginaire/generators/spade.py:394-395"""
    return x.repeat_interleave(2, dim=2).repeat_interleave(2, dim=3)


def _normalize(v, eps=1e-12):
    return v / max(float(v.norm()), eps)


def effective_weight(sd, prefix, training):
    """Weight of a (possibly spectrally normalised) conv / linear layer.
    ref: torch.nn.utils.spectral_norm as used by imaginaire/layers/weight_norm.py:85
    (dim 0, one power iteration per training-mode forward, eps 1e-12).  In training mode
    the updated u, v are written back into ``sd`` like the module's buffers."""
    if prefix + 'weight_orig' not in sd:
        return sd[prefix + 'weight']
    w = sd[prefix + 'weight_orig']
    mat = w.reshape(w.shape[0], -1)
    u, v = sd[prefix + 'weight_u'], sd[prefix + 'weight_v']
    if training:
        with torch.no_grad():   # u, v are constants for autograd, as in torch's implementation
            v = _normalize(torch.mv(mat.t(), u))
            u = _normalize(torch.mv(mat, v))
        sd[prefix + 'weight_u'], sd[prefix + 'weight_v'] = u, v
    sigma = torch.dot(u, torch.mv(mat, v))
    return w / sigma


def conv(sd, prefix, x, stride, padding, training, store=True):
    """ref: imaginaire/layers/conv.py:93-104 (nn.Conv2d, zero padding)"""
    y = F.conv2d(_q(x), _q(effective_weight(sd, prefix, training)), sd.get(prefix + 'bias'), stride=stride,
                 padding=padding)
    return _s(y) if store else y


def linear(sd, prefix, x, training):
    return F.linear(x, effective_weight(sd, prefix, training), sd.get(prefix + 'bias'))


def leaky_relu(x):
    """LeakyReLU(0.2). ref: imaginaire/layers/nonlinearity.py:22"""
    return torch.where(x > 0, x, 0.2 * x)


def inner_norm(sd, prefix, x, kind, training, momentum=0.1):
    """The norm inside SPADE / AdaptiveNorm, built by get_activation_norm_layer.
    ref: imaginaire/layers/activation_norm.py:396-410.  Biased variance normalises, the
    unbiased one updates running_var; instance norm always uses instance statistics."""
    if kind in ('none', ''):
        return x
    w, b = sd.get(prefix + 'weight'), sd.get(prefix + 'bias')
    if kind == 'instance':
        mean = x.mean((2, 3), keepdim=True)
        var = x.var((2, 3), unbiased=False, keepdim=True)
    elif kind in ('batch', 'sync_batch'):
        if training:
            mean = x.mean((0, 2, 3), keepdim=True)
            var = x.var((0, 2, 3), unbiased=False, keepdim=True)
            n = x.numel() / x.shape[1]
            rm, rv = sd[prefix + 'running_mean'], sd[prefix + 'running_var']
            sd[prefix + 'running_mean'] = (1 - momentum) * rm + momentum * mean.flatten()
            sd[prefix + 'running_var'] = (1 - momentum) * rv + momentum * var.flatten() * (n / (n - 1))
            if prefix + 'num_batches_tracked' in sd:
                sd[prefix + 'num_batches_tracked'] = sd[prefix + 'num_batches_tracked'] + 1
        else:
            mean = sd[prefix + 'running_mean'].view(1, -1, 1, 1)
            var = sd[prefix + 'running_var'].view(1, -1, 1, 1)
    else:
        raise ValueError(kind)
    y = (x - mean) / torch.sqrt(var + EPS_BN)
    if w is not None:
        y = y * w.view(1, -1, 1, 1) + b.view(1, -1, 1, 1)
    return y


# -------------------------------------------------------------------- SPADE

def spatially_adaptive_norm(sd, prefix, x, cond_inputs, p, training):
    """ref: SpatiallyAdaptiveNorm.forward, imaginaire/layers/activation_norm.py:211-234.
    p: dict(kernel_size, separate_projection, num_filters, norm_kind)."""
    pad = p['kernel_size'] // 2
    out = inner_norm(sd, prefix + 'norm.', x, p['norm_kind'], training)
    for i, cond in enumerate(cond_inputs):
        if cond is None:
            continue
        label = nearest_resize(cond, x.shape[2:])
        mlp = prefix + 'mlps.%d.' % i
        if p['separate_projection']:
            hidden = label
            if p['num_filters'] > 0:
                hidden = F.relu(conv(sd, mlp + '0.layers.conv.', label, 1, pad, training))
            gamma = conv(sd, prefix + 'gammas.%d.layers.conv.' % i, hidden, 1, pad, training, store=STORE_GAMMA_BETA)
            beta = conv(sd, prefix + 'betas.%d.layers.conv.' % i, hidden, 1, pad, training, store=STORE_GAMMA_BETA)
        else:
            hidden, k = label, 0
            if p['num_filters'] > 0:
                hidden = F.relu(conv(sd, mlp + '0.layers.conv.', label, 1, pad, training))
                k = 1
            gamma, beta = conv(sd, mlp + '%d.layers.conv.' % k, hidden, 1, pad, training,
                               store=STORE_GAMMA_BETA).chunk(2, dim=1)
        out = out * (1 + gamma) + beta
    return out


def adaptive_norm(sd, prefix, x, y, norm_kind, separate_projection, training):
    """ref: AdaptiveNorm.forward, imaginaire/layers/activation_norm.py:79-106"""
    if separate_projection:
        gamma = linear(sd, prefix + 'fc_gamma.layers.conv.', y, training)
        beta = linear(sd, prefix + 'fc_beta.layers.conv.', y, training)
    else:
        gamma, beta = linear(sd, prefix + 'fc.layers.conv.', y, training).chunk(2, 1)
    x = inner_norm(sd, prefix + 'norm.', x, norm_kind, training)
    return x * (1 + gamma[:, :, None, None]) + beta[:, :, None, None]


def spade_res_block(sd, prefix, x, seg, p, training, kernel_size=3):
    """Res2dBlock with order 'NACNAC', SPADE norms, LeakyReLU, learned 1x1 shortcut when the
    channel counts differ.  ref: _BaseResBlock, imaginaire/layers/residual.py:16-151."""
    pad = int(math.ceil((kernel_size - 1.0) / 2))
    dx = spatially_adaptive_norm(sd, prefix + 'conv_block_0.layers.norm.', x, [seg], p, training)
    dx = conv(sd, prefix + 'conv_block_0.layers.conv.', leaky_relu(dx), 1, pad, training)
    dx = spatially_adaptive_norm(sd, prefix + 'conv_block_1.layers.norm.', dx, [seg], p, training)
    # the residual sum is formed on the fp32 accumulators of this convolution and stored once
    dx = conv(sd, prefix + 'conv_block_1.layers.conv.', leaky_relu(dx), 1, pad, training, store=False)
    if (prefix + 'conv_block_s.layers.conv.weight_orig' in sd) or (prefix + 'conv_block_s.layers.conv.weight' in sd):
        xs = x
        if (prefix + 'conv_block_s.layers.norm.norm.weight' in sd) or \
                any(k.startswith(prefix + 'conv_block_s.layers.norm.') for k in sd):
            xs = spatially_adaptive_norm(sd, prefix + 'conv_block_s.layers.norm.', x, [seg], p, training)
        xs = conv(sd, prefix + 'conv_block_s.layers.conv.', xs, 1, 0, training)   # no activation (:98-108)
    else:
        xs = x
    return _s(xs + dx)


def style_encoder(sd, prefix, images, kernel_size, training, eps_noise):
    """ref: StyleEncoder.forward, imaginaire/generators/spade.py:535-563"""
    pad = int(math.ceil((kernel_size - 1.0) / 2))
    x = images
    if x.shape[2] != 256 or x.shape[3] != 256:
        x = F.interpolate(x, size=(256, 256), mode='bilinear')
    for i in range(1, 7):
        x = leaky_relu(conv(sd, prefix + 'layer%d.layers.conv.' % i, x, 2, pad, training))
    x = x.reshape(x.shape[0], -1)
    mu = linear(sd, prefix + 'fc_mu.layers.conv.', x, training)
    logvar = linear(sd, prefix + 'fc_var.layers.conv.', x, training)
    z = eps_noise * torch.exp(0.5 * logvar) + mu
    return mu, logvar, z


def posenc_grid():
    """ref: imaginaire/generators/spade.py:396-399"""
    xv, yv = torch.meshgrid([torch.arange(-1, 1.1, 2. / 15), torch.arange(-1, 1.1, 2. / 15)], indexing='ij')
    return torch.cat((xv.unsqueeze(0), yv.unsqueeze(0)), 0).unsqueeze(0)


def spade_generator(sd, seg, z, g, training, prefix='spade_generator.'):
    """ref: SPADEGenerator.forward, imaginaire/generators/spade.py:401-493.
    g: dict(kernel_size, small_side, use_style, use_posenc, global_norm, spade=dict(...))."""
    k = g['kernel_size']
    pad = int(math.ceil((k - 1.0) / 2))
    p = g['spade']
    base = {256: 16, 512: 32, 1024: 64}[g['small_side']]
    if g['use_style']:
        z = F.relu(linear(sd, prefix + 'fc_0.layers.conv.', z, training))
        z = F.relu(linear(sd, prefix + 'fc_1.layers.conv.', z, training))
    sy, sx = seg.shape[2] // base, seg.shape[3] // base
    in_seg = nearest_resize(seg, (sy, sx))
    if g['use_posenc']:
        in_xy = F.interpolate(posenc_grid(), size=[sy, sx], mode='bicubic')
        in_seg = torch.cat((in_seg, in_xy.expand(in_seg.shape[0], 2, sy, sx)), 1)

    def mid(name, x):
        if g['use_style']:
            h = adaptive_norm(sd, prefix + 'cbn_%s.layers.norm.' % name, x, z, g['global_norm'],
                              p['separate_projection'], training)
            return _trace('cbn_' + name, conv(sd, prefix + 'cbn_%s.layers.conv.' % name, leaky_relu(h), 1, pad, training))
        return _trace('conv_' + name, conv(sd, prefix + 'conv_%s.layers.conv.' % name, leaky_relu(x), 1, pad, training))

    def res(name, x):
        return _trace(name, spade_res_block(sd, prefix + name + '.', x, seg, p, training, k))

    def to_rgb(name, x):
        return conv(sd, prefix + name + '.layers.conv.', leaky_relu(x), 1, 2, training)

    x = _trace('head_0', leaky_relu(conv(sd, prefix + 'head_0.layers.conv.', in_seg, 1, pad, training)))
    x = mid('head_0', x)
    x = res('head_2', res('head_1', x))
    x = upsample2x(x)
    x = res('up_0b', mid('up_0a', res('up_0a', x)))
    x = upsample2x(x)
    x = res('up_1b', mid('up_1a', res('up_1a', x)))
    x = upsample2x(x)
    x = res('up_2b', mid('up_2a', res('up_2a', x)))
    x = upsample2x(x)
    if g['small_side'] == 256:
        return torch.tanh(to_rgb('conv_img256', x))
    x256 = upsample2x(to_rgb('conv_img256', x))
    x = res('up_3b', res('up_3a', x))
    x = upsample2x(x)
    x512 = to_rgb('conv_img512', x)
    if g['small_side'] == 512:
        return torch.tanh(x256 + x512)
    x256, x512 = upsample2x(x256), upsample2x(x512)
    x = res('up_4b', res('up_4a', x))
    x = upsample2x(x)
    return torch.tanh(x256 + x512 + to_rgb('conv_img1024', x))


def generator(sd, data, g, training, eps_noise=None):
    """ref: Generator.forward, imaginaire/generators/spade.py:131-166.  Returns a dict like
    the reference.  ``sd`` is copied first; the (training-mode) updated buffers come back
    under 'state_dict'."""
    sd = {k: v.clone() for k, v in sd.items()}
    mu = logvar = z = None
    if g['use_style']:
        mu, logvar, z = style_encoder(sd, 'style_encoder.', data['images'], g['style_kernel_size'], training,
                                      eps_noise)
    out = spade_generator(sd, data['label'], z, g, training)
    return {'fake_images': out, 'mu': mu, 'logvar': logvar, 'state_dict': sd}


def generator_spec_from_cfg(gen_cfg, crop_hw):
    """The few numbers the functional oracle needs from a generator config."""
    an = getattr(gen_cfg, 'activation_norm_params', None)
    get = (lambda k, d: getattr(an, k, d)) if an is not None else (lambda k, d: d)
    style_enc = getattr(gen_cfg, 'style_enc', None)
    return dict(
        kernel_size=getattr(gen_cfg, 'kernel_size', 3),
        small_side=min(crop_hw),
        use_style=getattr(gen_cfg, 'style_dims', None) is not None,
        use_posenc=getattr(gen_cfg, 'use_posenc_in_input_layer', True),
        global_norm=getattr(gen_cfg, 'global_adaptive_norm_type', 'sync_batch'),
        style_kernel_size=getattr(style_enc, 'kernel_size', 3) if style_enc is not None else 3,
        spade=dict(kernel_size=get('kernel_size', 3), separate_projection=get('separate_projection', False),
                   num_filters=get('num_filters', 128), norm_kind=get('activation_norm_type', 'sync_batch')))
