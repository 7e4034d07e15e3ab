"""Convolution / linear blocks with the reference's module API, running on the native
sm_100a kernels.

ref: imaginaire/layers/conv.py:14-135 (_BaseConvBlock), :138-191 (LinearBlock),
     :262-327 (Conv2dBlock)

Same constructor arguments, same ``layers`` ModuleDict (keys ``conv`` / ``noise`` /
``norm`` / ``nonlinearity`` in ``order``), same ``conditional`` attribute and the same
parameter names, so a reference ``state_dict`` loads with ``strict=True``.  What
differs is how ``forward`` executes: a 2-D block converts its input once to the
internal activation layout (NHWC bf16) and runs

  * 'C'  -> the tcgen05 implicit-GEMM convolution (bias, a following LeakyReLU/ReLU,
            a residual add and the next norm's statistics fused into its epilogue),
  * 'N'(+'A') -> one fused normalise * (1+gamma) + beta -> activation kernel,

instead of one cuDNN / ATen launch per operation.  There is no cuDNN dispatch and
no CPU path: CPU tensors raise.
"""
from types import SimpleNamespace

import torch
from torch import nn
from torch.nn import functional as F
from torch.nn.utils.spectral_norm import SpectralNorm

from .. import functional as NF
from .. import ops
from .misc import ApplyNoise


class Conv2d(nn.Conv2d):
    """nn.Conv2d as the parameter owner (identical init, keys and spectral-norm hooks);
    the computation is ops.conv2d_fprop."""

    def effective_weight(self):
        """Runs the weight-norm forward pre-hooks (spectral norm's power iteration and
        W / sigma) without calling the module, and returns the weight to convolve with."""
        for hook in self._forward_pre_hooks.values():
            hook(self, None)
        return self.weight

    def _check_supported(self):
        if self.groups != 1 or tuple(self.dilation) != (1, 1) or self.padding_mode != 'zeros':
            raise NotImplementedError('native Conv2d supports groups=1, dilation=1, zero padding only')
        if self.kernel_size[0] != self.kernel_size[1] or self.stride[0] != self.stride[1] or \
                self.padding[0] != self.padding[1] or self.stride[0] not in (1, 2):
            raise NotImplementedError('native Conv2d supports square kernels, stride 1 or 2, symmetric padding')

    def spectral_hook(self):
        """torch's SpectralNorm hook on `weight` (dim 0), if this layer has one."""
        for hook in self._forward_pre_hooks.values():
            if isinstance(hook, SpectralNorm) and hook.name == 'weight' and hook.dim == 0:
                return hook
        return None

    def weight_for_kernel(self):
        """(filter to pack, sn): for a spectrally normalised layer the filter is weight_orig and
        sn = (1/sigma, u, v) after this forward's power iteration (the division is folded into the
        repack kernel and the backward handles sigma); otherwise (effective weight, None)."""
        hook = self.spectral_hook()
        if hook is None:
            return self.effective_weight(), None
        ready = self.__dict__.pop('_sn_ready', None)
        if ready is not None:
            return self.weight_orig, ready      # computed for the whole network by prefetch_spectral_norms
        return self.weight_orig, NF.spectral_sigma(self.weight_orig, self.weight_u, self.weight_v, self.training,
                                                   hook.n_power_iterations, hook.eps)

    def native_forward(self, x, act=ops.ACT_NONE, residual=None, stats_mode=None):
        self._check_supported()
        weight, sn = self.weight_for_kernel()
        return NF.conv2d(x, weight, self.bias, residual, self.stride[0], self.padding[0], act, stats_mode, sn)

    def forward(self, x):
        return self.native_forward(NF.to_nhwc(x))


def prefetch_spectral_norms(root):
    """Runs this forward's spectral-norm power iteration and sigma of EVERY spectrally normalised
    convolution under `root` in one batched call (imx_spectral_norm_sigma_batched_f32: five
    launches instead of four per layer).  The iteration depends on the filters only, so doing all
    layers up front is the same computation as the reference's per-layer hook; each layer picks
    its (1/sigma, u, v) up in weight_for_kernel().  Call it only for modules that WILL run in
    this forward: like the hook, every call advances u and v once."""
    layers = root.__dict__.get('_sn_layers')
    if layers is None:      # the module tree is static: walk it once
        layers = [m for m in root.modules() if isinstance(m, Conv2d) and m.spectral_hook() is not None]
        root.__dict__['_sn_layers'] = layers
    if not layers or not layers[0].weight_orig.is_cuda:
        return
    hooks = [m.spectral_hook() for m in layers]
    if any(h.n_power_iterations != 1 or h.eps != hooks[0].eps for h in hooks):
        return                      # unusual settings: every layer falls back to its own call
    key = tuple((m.weight_orig.data_ptr(), m.weight_u.data_ptr(), m.weight_v.data_ptr()) for m in layers)
    plan = root.__dict__.get('_sn_plan')
    if plan is None or plan['key'] != key:
        dev = layers[0].weight_orig.device
        rows, t_off, out_off, spans = [], 0, 0, []
        for m in layers:
            r = m.weight_orig.shape[0]
            c = m.weight_orig.numel() // r
            rows.append([m.weight_orig.data_ptr(), m.weight_u.data_ptr(), m.weight_v.data_ptr(), t_off, 0, out_off, r, c])
            spans.append((out_off, r, c))
            t_off += c
            out_off += 1 + r + c
        s_off = t_off
        for row in rows:
            row[4] = s_off
            s_off += row[6]
        plan = dict(key=key, descs=torch.tensor(rows, dtype=torch.int64).to(dev), t_floats=t_off,
                    scratch=torch.empty((s_off,), dtype=torch.float32, device=dev), out_floats=out_off, spans=spans,
                    max_r=max(r[6] for r in rows), max_c=max(r[7] for r in rows))
        root.__dict__['_sn_plan'] = plan
    training = layers[0].training
    with torch.no_grad():
        if training:
            plan['scratch'][:plan['t_floats']].zero_()
        out = ops.spectral_norm_sigma_batched(plan['descs'], len(layers), plan['max_r'], plan['max_c'], plan['scratch'],
                                              plan['out_floats'], 1 if training else 0, hooks[0].eps)
    for m, (off, r, c) in zip(layers, plan['spans']):
        m.__dict__['_sn_ready'] = (out[off:off + 1], out[off + 1:off + 1 + r], out[off + 1 + r:off + 1 + r + c])


_FUSABLE_ACT = {nn.LeakyReLU: ops.ACT_LRELU, nn.ReLU: ops.ACT_RELU}


def _act_code(layer):
    """ACT_* code if the activation module can be folded into a native kernel, else None."""
    code = _FUSABLE_ACT.get(type(layer))
    if code == ops.ACT_LRELU and abs(layer.negative_slope - 0.2) > 1e-12:
        return None
    return code


class _BaseConvBlock(nn.Module):
    def __init__(self, in_channels, out_channels, kernel_size, stride,
                 padding, dilation, groups, bias, padding_mode,
                 weight_norm_type, weight_norm_params,
                 activation_norm_type, activation_norm_params,
                 nonlinearity, inplace_nonlinearity,
                 apply_noise, order, input_dim):
        super().__init__()
        from .activation_norm import get_activation_norm_layer
        from .nonlinearity import get_nonlinearity_layer
        from .weight_norm import get_weight_norm_layer
        self.weight_norm_type = weight_norm_type
        self.input_dim = input_dim
        self._out_channels = out_channels

        # Construction order (conv -> noise -> norm -> nonlinearity) follows the reference so
        # that a seeded build draws the same random numbers (conv.py:32-57).
        wn_params = vars(weight_norm_params) if weight_norm_params is not None else {}
        conv_layer = get_weight_norm_layer(weight_norm_type, **wn_params)(
            self._get_conv_layer(in_channels, out_channels, kernel_size, stride, padding, dilation,
                                 groups, bias, padding_mode, input_dim))
        noise_layer = ApplyNoise() if apply_noise else None
        conv_first = order.find('C') < order.find('N')
        an_params = vars(activation_norm_params) if activation_norm_params is not None else {}
        norm_layer = get_activation_norm_layer(out_channels if conv_first else in_channels,
                                               activation_norm_type, input_dim, **an_params)
        act_layer = get_nonlinearity_layer(nonlinearity, inplace=inplace_nonlinearity)

        self.layers = nn.ModuleDict()
        for op in order:
            name, layer = {'C': ('conv', conv_layer), 'N': ('norm', norm_layer),
                           'A': ('nonlinearity', act_layer)}[op]
            if layer is None:
                continue
            self.layers[name] = layer
            if op == 'C' and noise_layer is not None:
                self.layers['noise'] = noise_layer

        self.conditional = getattr(conv_layer, 'conditional', False) or \
            getattr(norm_layer, 'conditional', False)

    def _get_conv_layer(self, in_channels, out_channels, kernel_size, stride, padding, dilation,
                        groups, bias, padding_mode, input_dim):
        if input_dim == 0:
            return nn.Linear(in_channels, out_channels, bias)
        if input_dim == 2:
            return Conv2d(in_channels, out_channels, kernel_size, stride, padding, dilation, groups,
                          bias, padding_mode)
        raise NotImplementedError('Conv%dd blocks are outside the SPADE hot path' % input_dim)

    # ---------------------------------------------------------------- forward
    def forward(self, x, *cond_inputs, **kw_cond_inputs):
        if self.input_dim == 2:
            y = self.forward_native(x, cond_inputs, kw_cond_inputs)
            # internal tensors carry channels padded to a multiple of 8 (zeros); callers of the
            # public API see the logical channel count
            return y if y.shape[1] == self._out_channels else y[:, :self._out_channels]
        # Linear blocks (style code MLPs): tiny GEMMs, plain torch (cuBLAS) calls.
        for layer in self.layers.values():
            if getattr(layer, 'conditional', False):
                x = layer(x, *cond_inputs, **kw_cond_inputs)
            else:
                x = layer(x)
        return x

    def forward_native(self, x, cond_inputs=(), kw_cond_inputs=None, residual=None, stats_mode=None,
                       post_act=ops.ACT_NONE, skip_first_act=False):
        """2-D forward on the native kernels.

        post_act: activation of the CONSUMER block applied in the epilogue of this block's last
            convolution (after bias and residual); the consumer is then run with skip_first_act=True.
            Only honoured when the convolution is the block's last operation and has no activation
            of its own (callers check with ``ends_with_plain_conv``).
        skip_first_act: the block's leading activation was already applied by the producer.
        residual: NHWC bf16 tensor added in the epilogue of the block's convolution.
        stats_mode: 'batch' / 'instance' -> the convolution's epilogue also produces the
            statistics that the next normalisation layer needs (only honoured when the
            convolution is the last operation of the block).
        """
        kw_cond_inputs = kw_cond_inputs or {}
        x = NF.to_nhwc(x)
        items = list(self.layers.items())
        i = 0
        while i < len(items):
            name, layer = items[i]
            nxt = items[i + 1] if i + 1 < len(items) else None
            fused_act = ops.ACT_NONE
            if name in ('conv', 'norm') and nxt is not None and nxt[0] == 'nonlinearity' and \
                    _act_code(nxt[1]) is not None:
                fused_act = _act_code(nxt[1])
            if name == 'conv':
                last = (i + (2 if fused_act else 1)) >= len(items)
                conv_act = fused_act
                if last and fused_act == ops.ACT_NONE and post_act != ops.ACT_NONE and stats_mode is None:
                    conv_act, post_act = post_act, ops.ACT_NONE      # the consumer's activation rides in this epilogue
                x = layer.native_forward(x, act=conv_act, residual=residual if last else None,
                                         stats_mode=stats_mode if last else None)
                if last:
                    residual = None
            elif name == 'norm':
                from .activation_norm import apply_norm_layer
                x = apply_norm_layer(layer, x, cond_inputs, kw_cond_inputs, fused_act)
            elif name == 'nonlinearity':
                if not (skip_first_act and i == 0):
                    x = self._standalone_act(layer, x)
            else:   # noise
                x = layer(x)
            i += 2 if fused_act else 1
        if residual is not None:
            x = x + residual
        if post_act != ops.ACT_NONE:     # could not ride in an epilogue: apply it here
            x = F.leaky_relu(x, 0.2) if post_act == ops.ACT_LRELU else F.relu(x)
        return x

    def ends_with_plain_conv(self):
        """True when the block's last operation is its convolution (so a consumer's activation can
        ride in that convolution's epilogue)."""
        return len(self.layers) > 0 and list(self.layers.keys())[-1] == 'conv'

    def leading_act_code(self):
        """Kernel code of the block's leading activation, or None if it does not start with a fusable one."""
        if len(self.layers) == 0:
            return None
        name, layer = next(iter(self.layers.items()))
        return _act_code(layer) if name == 'nonlinearity' else None

    @staticmethod
    def _standalone_act(layer, x):
        # An activation with no producer in this block (order 'A..'): elementwise torch op
        # on the bf16 NHWC tensor (layout preserved).
        if isinstance(layer, nn.LeakyReLU):
            return F.leaky_relu(x, layer.negative_slope)
        if isinstance(layer, nn.ReLU):
            return F.relu(x)
        return layer(x)

    def __repr__(self):
        lines = []
        for name, layer in self.layers.items():
            text = repr(layer)
            if name == 'conv' and self.weight_norm_type not in ('none', ''):
                text = text[:-1] + ', weight_norm={})'.format(self.weight_norm_type)
            lines.append('  ' + text.replace('\n', '\n  '))
        return self._get_name() + '(\n' + '\n'.join(lines) + '\n)'


class LinearBlock(_BaseConvBlock):
    """ref: imaginaire/layers/conv.py:138-191"""

    def __init__(self, in_features, out_features, bias=True,
                 weight_norm_type='none', weight_norm_params=None,
                 activation_norm_type='none', activation_norm_params=None,
                 nonlinearity='none', inplace_nonlinearity=False,
                 apply_noise=False, order='CNA'):
        super().__init__(in_features, out_features, None, None, None, None, None, bias, None,
                         weight_norm_type, weight_norm_params,
                         activation_norm_type, activation_norm_params,
                         nonlinearity, inplace_nonlinearity, apply_noise, order, 0)


class Conv2dBlock(_BaseConvBlock):
    """ref: imaginaire/layers/conv.py:262-327"""

    def __init__(self, in_channels, out_channels, kernel_size, stride=1,
                 padding=0, dilation=1, groups=1, bias=True,
                 padding_mode='zeros',
                 weight_norm_type='none', weight_norm_params=None,
                 activation_norm_type='none', activation_norm_params=None,
                 nonlinearity='none', inplace_nonlinearity=False,
                 apply_noise=False, order='CNA'):
        super().__init__(in_channels, out_channels, kernel_size, stride, padding, dilation, groups, bias,
                         padding_mode, weight_norm_type, weight_norm_params,
                         activation_norm_type, activation_norm_params,
                         nonlinearity, inplace_nonlinearity, apply_noise, order, 2)
