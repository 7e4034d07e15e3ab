"""Activation normalisation layers of the SPADE hot path on the native kernels.

ref: imaginaire/layers/activation_norm.py:22-106 (AdaptiveNorm),
     :109-234 (SpatiallyAdaptiveNorm), :377-432 (get_activation_norm_layer)

The inner norm (``self.norm``) stays a torch.nn module so that parameters, running
statistics and state-dict keys are exactly the reference's; its ``forward`` is never
called.  Statistics come from the producing convolution's epilogue or the standalone
statistics kernel, SyncBatchNorm exchanges [sum, sumsq, count] with one NCCL
all-reduce, and normalise + affine + (1+gamma)*x+beta + activation is one kernel.
Under ``torch.no_grad()`` (inference, and the generator pass of the discriminator
update) the gamma/beta projection convolution applies the modulation in its own
epilogue, so gamma and beta are never written to HBM.
"""
from types import SimpleNamespace

import torch
from torch import nn
from torch.nn import functional as F

from .. import functional as NF
from .. import ops
from .conv import Conv2dBlock, LinearBlock


# ------------------------------------------------------------- inner norms

def _frozen_identity_affine(weight, bias):
    """True for a frozen weight / bias pair that is exactly (1, 0).

    The reference builds SyncBatchNorm with affine=True whatever the config says and only switches
    ``requires_grad`` off for ``affine: false`` (imaginaire/layers/activation_norm.py:403-410), so
    SPADE's "parameter-free" inner norm carries a weight of ones and a bias of zeros that never
    change.  Values are checked, not assumed (a loaded checkpoint may hold anything); the check
    reads the device once per tensor version."""
    if weight is None or bias is None or weight.requires_grad or bias.requires_grad or not weight.is_cuda:
        return False
    if getattr(weight, '_imx_trained', False) or getattr(bias, '_imx_trained', False):
        # the optimizer has moved them (the reference's trainer unfreezes every generator parameter for the
        # generator update, trainers/base.py:602): no longer (1, 0), and no device read-back per step to find out
        return False
    hit = getattr(weight, '_imx_identity_affine', None)     # cached on the tensor: dies with it
    if hit is None or hit[0] != weight._version or hit[1] != bias._version or hit[2] is not bias:
        with torch.no_grad():
            ident = bool(((weight == 1).all() & (bias == 0).all()).item())
        hit = (weight._version, bias._version, bias, ident)
        weight._imx_identity_affine = hit
    return hit[3]


def norm_state(norm):
    """NormState view of a torch.nn normalisation module (None -> identity)."""
    if norm is None:
        return None
    if isinstance(norm, nn.SyncBatchNorm):
        kind = 'sync_batch'
    elif isinstance(norm, nn.BatchNorm2d):
        kind = 'batch'
    elif isinstance(norm, nn.InstanceNorm2d):
        kind = 'instance'
        if norm.track_running_stats:
            raise NotImplementedError('InstanceNorm2d(track_running_stats=True) is not supported')
    else:
        raise NotImplementedError('inner norm %s is outside the SPADE hot path' % type(norm).__name__)
    weight, bias = getattr(norm, 'weight', None), getattr(norm, 'bias', None)
    if _frozen_identity_affine(weight, bias):
        weight = bias = None      # the kernels' affine-free instantiation (fewer live registers)
    return NF.NormState(
        kind, weight, bias,
        getattr(norm, 'running_mean', None), getattr(norm, 'running_var', None),
        norm.momentum, norm.eps, norm.training, norm.track_running_stats,
        getattr(norm, 'num_batches_tracked', None), getattr(norm, 'process_group', None))


class _IdentityStats:
    """mean 0 / rstd 1 'statistics' used to apply a further (1+gamma)*y+beta modulation
    (multi-SPADE) or a modulation without inner norm through the same fused kernel."""
    _cache = {}

    @classmethod
    def get(cls, c, device):
        key = (c, device)
        if key not in cls._cache:
            cls._cache[key] = (torch.zeros(c, device=device), torch.ones(c, device=device))
        return cls._cache[key]


def _modulate_only(x, gamma, beta, act):
    mean, rstd = _IdentityStats.get(x.shape[1], x.device)
    return NF._SpadeNormAct.apply(x, None, None, gamma, beta, mean, rstd, 1, False, act, False, None)


def apply_norm_layer(layer, x, cond_inputs, kw_cond_inputs, act):
    """Runs one 'N' (+ fused 'A') step of a conv block on an NHWC bf16 tensor."""
    if isinstance(layer, (SpatiallyAdaptiveNorm, AdaptiveNorm)):
        return layer.forward_fused(x, *cond_inputs, act=act, **kw_cond_inputs)
    ns = norm_state(layer)   # plain batch / sync-batch / instance norm
    return NF.spade_norm_act(x, ns, None, None, act)


# ---------------------------------------------------------------- adaptive

class AdaptiveNorm(nn.Module):
    """norm(x) * (1 + gamma(y)) + beta(y) with per-sample gamma / beta from a linear
    projection of the style code.  ref: activation_norm.py:22-106"""

    def __init__(self, num_features, cond_dims, weight_norm_type='',
                 projection=True, separate_projection=False, input_dim=2,
                 activation_norm_type='instance', activation_norm_params=None):
        super().__init__()
        if input_dim != 2:
            raise NotImplementedError('AdaptiveNorm: only 2-D inputs are on the hot path')
        self.projection = projection
        self.separate_projection = separate_projection
        if activation_norm_params is None:
            activation_norm_params = SimpleNamespace(affine=False)
        self.norm = get_activation_norm_layer(num_features, activation_norm_type, input_dim,
                                              **vars(activation_norm_params))
        if projection:
            if separate_projection:
                self.fc_gamma = LinearBlock(cond_dims, num_features, weight_norm_type=weight_norm_type)
                self.fc_beta = LinearBlock(cond_dims, num_features, weight_norm_type=weight_norm_type)
            else:
                self.fc = LinearBlock(cond_dims, num_features * 2, weight_norm_type=weight_norm_type)
        self.conditional = True

    def _gamma_beta(self, y):
        if self.projection:
            if self.separate_projection:
                return self.fc_gamma(y), self.fc_beta(y)
            y = self.fc(y)
        gamma, beta = y.chunk(2, 1)
        return gamma, beta

    def forward_fused(self, x, y, act=ops.ACT_NONE, **kwargs):
        x = NF.to_nhwc(x)
        gamma, beta = self._gamma_beta(y)
        gamma = gamma.float().contiguous()
        beta = beta.float().contiguous()
        ns = norm_state(self.norm)
        if ns is None:
            return _modulate_only(x, gamma, beta, act)
        return NF.spade_norm_act(x, ns, gamma, beta, act)

    def forward(self, x, y, **kwargs):
        return self.forward_fused(x, y, act=ops.ACT_NONE, **kwargs)


# ------------------------------------------------------------------- SPADE

class SpatiallyAdaptiveNorm(nn.Module):
    """Spatially-adaptive normalisation.  ref: activation_norm.py:109-234"""

    def __init__(self, num_features, cond_dims, num_filters=128, kernel_size=3,
                 weight_norm_type='', separate_projection=False,
                 activation_norm_type='sync_batch', activation_norm_params=None,
                 partial=False):
        super().__init__()
        if activation_norm_params is None:
            activation_norm_params = SimpleNamespace(affine=False)
        padding = kernel_size // 2
        self.separate_projection = separate_projection
        self.num_features = num_features
        self.kernel_size = kernel_size
        self.mlps = nn.ModuleList()
        self.gammas = nn.ModuleList()
        self.betas = nn.ModuleList()

        if type(cond_dims) != list:
            cond_dims = [cond_dims]
        if not isinstance(num_filters, list):
            num_filters = [num_filters] * len(cond_dims)
        else:
            assert len(num_filters) >= len(cond_dims)
        if isinstance(partial, list):
            assert len(partial) >= len(cond_dims)
            partial = any(partial)
        if partial:
            raise NotImplementedError('partial-convolution SPADE is outside the hot path')

        # projections first, inner norm last: the reference's construction (and RNG) order
        for i, cond_dim in enumerate(cond_dims):
            mlp = []
            if num_filters[i] > 0:
                mlp.append(Conv2dBlock(cond_dim, num_filters[i], kernel_size, padding=padding,
                                       weight_norm_type=weight_norm_type, nonlinearity='relu'))
            mlp_ch = cond_dim if num_filters[i] == 0 else num_filters[i]
            if separate_projection:
                self.mlps.append(nn.Sequential(*mlp))
                self.gammas.append(Conv2dBlock(mlp_ch, num_features, kernel_size, padding=padding,
                                               weight_norm_type=weight_norm_type))
                self.betas.append(Conv2dBlock(mlp_ch, num_features, kernel_size, padding=padding,
                                              weight_norm_type=weight_norm_type))
            else:
                mlp.append(Conv2dBlock(mlp_ch, num_features * 2, kernel_size, padding=padding,
                                       weight_norm_type=weight_norm_type))
                self.mlps.append(nn.Sequential(*mlp))

        self.norm = get_activation_norm_layer(num_features, activation_norm_type, 2,
                                              **vars(activation_norm_params))
        self.conditional = True

    # -- conditioning map at the feature resolution, once per resolution per forward ------
    @staticmethod
    def _resized_cond(cond, h, w):
        if ops.is_nhwc_bf16(cond):
            if cond.shape[2] == h and cond.shape[3] == w:
                return cond
            # feature maps of an embedding network used below their own resolution (vid2vid's
            # coarsest SPADE layers): nearest resize in place of the layout (differentiable)
            return F.interpolate(cond, size=(h, w), mode='nearest')
        if cond.requires_grad:
            # learned conditioning maps (vid2vid label embeddings): keep it differentiable
            return NF.to_nhwc(F.interpolate(cond.float(), size=(h, w), mode='nearest'))
        # kept on the label tensor (dies with the batch); a caller that refills the same tensor in place
        # (static input buffer, prefetcher copy_) advances its version counter and the maps are rebuilt
        cache = cond.__dict__.get('_imx_resized')
        if cache is None or cache[0] != cond._version:
            cache = (cond._version, {})
            cond.__dict__['_imx_resized'] = cache
        key = (h, w)
        if key not in cache[1]:
            cache[1][key] = ops.nearest_resize_to_nhwc(cond, h, w)
        return cache[1][key]

    def _projection(self, i):
        """(hidden block or None, gamma conv, beta conv or None) of conditioning input i."""
        mlp = self.mlps[i]
        if self.separate_projection:
            hidden = mlp[0] if len(mlp) > 0 else None
            return hidden, self.gammas[i].layers['conv'], self.betas[i].layers['conv']
        hidden = mlp[0] if len(mlp) > 1 else None
        return hidden, mlp[-1].layers['conv'], None

    def _gamma_beta_weights(self, i):
        """Filters / biases of gamma and beta: (wg, wb, bg, bb, sng, snb, conv).  Separate
        projections go through the fused spectral-norm path (weight_orig + 1/sigma); a joint
        projection is split into its two halves of the effective weight."""
        _, conv_g, conv_b = self._projection(i)
        c = self.num_features
        if conv_b is not None:
            wg, sng = conv_g.weight_for_kernel()
            wb, snb = conv_b.weight_for_kernel()
            return wg, wb, conv_g.bias, conv_b.bias, sng, snb, conv_g
        w = conv_g.effective_weight()
        bias = conv_g.bias
        return (w[:c], w[c:], (None if bias is None else bias[:c]), (None if bias is None else bias[c:]),
                None, None, conv_g)

    def forward_fused(self, x, *cond_inputs, act=ops.ACT_NONE, **kwargs):
        x = NF.to_nhwc(x)
        n, c, h, w = x.shape
        active = [i for i in range(len(cond_inputs)) if cond_inputs[i] is not None]
        ns = norm_state(self.norm)
        if not active:
            if ns is None:
                return x if act == ops.ACT_NONE else _modulate_only(x, None, None, act)
            return NF.spade_norm_act(x, ns, None, None, act)
        out = x
        for k, i in enumerate(active):
            is_last = k == len(active) - 1
            step_act = act if is_last else ops.ACT_NONE
            label = self._resized_cond(cond_inputs[i], h, w)
            hidden_block, _, _ = self._projection(i)
            hidden = hidden_block(label) if hidden_block is not None else NF.to_nhwc(label)
            wg, wb, bg, bb, sng, snb, conv = self._gamma_beta_weights(i)
            conv._check_supported()
            pad = conv.padding[0]
            fuse = (not torch.is_grad_enabled()) and conv.stride[0] == 1 and \
                2 * pad == conv.kernel_size[0] - 1 and c % 32 == 0
            step_ns = ns if k == 0 else None
            if fuse:
                if step_ns is None:
                    mean, rstd = _IdentityStats.get(c, x.device)
                    step_ns = NF.NormState('batch', None, None, mean, torch.ones_like(rstd) - 1e-5, 0.1, 1e-5,
                                           False, True)
                out = NF.spade_gb_fused(hidden, wg, wb, bg, bb, out, step_ns, step_act, pad,
                                        None if sng is None else sng[0], None if snb is None else snb[0])
            else:
                gamma = NF.conv2d(hidden, wg, bg, None, 1, pad, sn=sng)
                beta = NF.conv2d(hidden, wb, bb, None, 1, pad, sn=snb)
                if step_ns is None:
                    out = _modulate_only(out, gamma, beta, step_act)
                else:
                    out = NF.spade_norm_act(out, step_ns, gamma, beta, step_act)
        return out

    def forward(self, x, *cond_inputs, **kwargs):
        return self.forward_fused(x, *cond_inputs, act=ops.ACT_NONE, **kwargs)


# ----------------------------------------------------------------- factory

def get_activation_norm_layer(num_features, norm_type, input_dim, **norm_params):
    """ref: imaginaire/layers/activation_norm.py:377-432"""
    input_dim = max(input_dim, 1)
    if norm_type in ('none', ''):
        return None
    if norm_type == 'batch':
        return getattr(nn, 'BatchNorm%dd' % input_dim)(num_features, **norm_params)
    if norm_type == 'instance':
        affine = norm_params.pop('affine', True)
        return getattr(nn, 'InstanceNorm%dd' % input_dim)(num_features, affine=affine, **norm_params)
    if norm_type == 'sync_batch':
        # always built with affine=True; `affine` only decides whether they train (:403-410)
        affine = norm_params.pop('affine', True)
        layer = nn.SyncBatchNorm(num_features, affine=True, **norm_params)
        layer.weight.requires_grad = affine
        layer.bias.requires_grad = affine
        return layer
    if norm_type == 'adaptive':
        return AdaptiveNorm(num_features, **norm_params)
    if norm_type == 'spatially_adaptive':
        if input_dim != 2:
            raise ValueError('Spatially adaptive normalization layers only supports 2D input')
        return SpatiallyAdaptiveNorm(num_features, **norm_params)
    if norm_type in ('layer', 'layer_2d', 'group', 'hyper_spatially_adaptive'):
        raise NotImplementedError("activation norm '%s' is outside the SPADE hot path "
                                  '(SURVEY.md section 2.1 #1)' % norm_type)
    raise ValueError('Activation norm layer %s is not recognized' % norm_type)
