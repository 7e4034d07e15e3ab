"""Residual blocks with the reference's module API on the native kernels.

ref: imaginaire/layers/residual.py:16-151 (_BaseResBlock), :305-382 (Res2dBlock)

Same constructor, same attributes (``conv_block_0`` / ``conv_block_1`` /
``conv_block_s``, ``learn_shortcut``, ``conditional``).  Execution differences:
the shortcut is computed first and added inside the epilogue of conv_block_1's
convolution (no separate add kernel), conv_block_0's convolution emits the
statistics conv_block_1's norm needs, and - when the parent set ``emit_stats`` -
the block's last convolution emits the statistics of the block output for the next
block's norm.  The two norms that see the block input (conv_block_0 and
conv_block_s) share one statistics pass: it is attached to the input tensor.
"""
import torch
from torch import nn
from torch.utils.checkpoint import checkpoint

from .. import functional as NF
from .. import ops
from .activation_norm import SpatiallyAdaptiveNorm, AdaptiveNorm, norm_state
from .conv import Conv2dBlock, LinearBlock


def _inner_stats_mode(block):
    """'batch' / 'instance' if `block` starts with a norm whose statistics a producing
    convolution can emit in its epilogue, else None."""
    layers = list(block.layers.items())
    if not layers or layers[0][0] != 'norm':
        return None
    norm = layers[0][1]
    inner = norm.norm if isinstance(norm, (SpatiallyAdaptiveNorm, AdaptiveNorm)) else norm
    ns = norm_state(inner) if inner is not None else None
    if ns is None or not ns.use_batch_stats:
        return None
    return 'instance' if ns.per_instance else 'batch'


class _BaseResBlock(nn.Module):
    def __init__(self, in_channels, out_channels, kernel_size,
                 padding, dilation, groups, bias, padding_mode,
                 weight_norm_type, weight_norm_params,
                 activation_norm_type, activation_norm_params,
                 skip_activation_norm, skip_nonlinearity,
                 nonlinearity, inplace_nonlinearity, apply_noise,
                 hidden_channels_equal_out_channels,
                 order, block, learn_shortcut):
        super().__init__()
        if order == 'pre_act':
            order = 'NACNAC'
        if isinstance(bias, bool):
            biases = [bias] * 3
        elif isinstance(bias, list) and len(bias) == 3:
            biases = bias   # conv_block_0, conv_block_1, conv_block_s
        else:
            raise ValueError('Bias must be a bool or a list of 3 bools.')
        if not 5 <= len(order) <= 6:
            raise ValueError('order must be either 5 or 6 characters')
        self.learn_shortcut = (in_channels != out_channels) or learn_shortcut
        hidden_channels = out_channels if hidden_channels_equal_out_channels \
            else min(in_channels, out_channels)

        main, skip = {}, {}
        if block != LinearBlock:
            base = dict(stride=1, dilation=dilation, groups=groups, padding_mode=padding_mode)
            main = dict(base, kernel_size=kernel_size, padding=padding,
                        activation_norm_type=activation_norm_type,
                        activation_norm_params=activation_norm_params)
            skip = dict(base, kernel_size=1)
            if skip_activation_norm:
                skip.update(activation_norm_type=activation_norm_type,
                            activation_norm_params=activation_norm_params)
        common = dict(weight_norm_type=weight_norm_type, weight_norm_params=weight_norm_params,
                      apply_noise=apply_noise)

        no_norm = activation_norm_type in ('', 'none')
        first_inplace = False if (order.find('A') < order.find('C') and no_norm) else inplace_nonlinearity
        self.conv_block_0 = block(in_channels, hidden_channels, bias=biases[0],
                                  nonlinearity=nonlinearity, order=order[0:3],
                                  inplace_nonlinearity=first_inplace, **main, **common)
        self.conv_block_1 = block(hidden_channels, out_channels, bias=biases[1],
                                  nonlinearity=nonlinearity, order=order[3:],
                                  inplace_nonlinearity=inplace_nonlinearity, **main, **common)
        if self.learn_shortcut:
            self.conv_block_s = block(in_channels, out_channels, bias=biases[2],
                                      nonlinearity=nonlinearity if skip_nonlinearity else '',
                                      order=order[0:3], **skip, **common)
        self.conditional = getattr(self.conv_block_0, 'conditional', False) or \
            getattr(self.conv_block_1, 'conditional', False)
        # set by the parent network: statistics mode of whatever normalises this block's output
        self.emit_stats = None
        self._native = block is Conv2dBlock

    def conv_blocks(self, x, *cond_inputs, **kw_cond_inputs):
        dx = self.conv_block_0(x, *cond_inputs, **kw_cond_inputs)
        return self.conv_block_1(dx, *cond_inputs, **kw_cond_inputs)

    def _forward_native(self, x, cond_inputs, kw_cond_inputs):
        x = NF.to_nhwc(x)
        if getattr(x, '_imx_stats', None) is None and _inner_stats_mode(self.conv_block_0) is not None \
                and self.learn_shortcut and _inner_stats_mode(self.conv_block_s) is not None:
            # conv_block_0.norm and conv_block_s.norm normalise the same tensor: one pass
            per_instance = _inner_stats_mode(self.conv_block_0) == 'instance'
            n, c, h, w = x.shape
            x._imx_stats = (ops.bn_stats(x, per_instance), per_instance, h * w if per_instance else n * h * w)
        shortcut = self.conv_block_s.forward_native(x, cond_inputs, kw_cond_inputs) \
            if self.learn_shortcut else x
        mid_stats = _inner_stats_mode(self.conv_block_1)
        dx = self.conv_block_0.forward_native(x, cond_inputs, kw_cond_inputs, stats_mode=mid_stats)
        post_act, self.post_act = getattr(self, 'post_act', ops.ACT_NONE), ops.ACT_NONE   # one-shot, set by the caller
        return self.conv_block_1.forward_native(dx, cond_inputs, kw_cond_inputs, residual=shortcut,
                                                stats_mode=self.emit_stats, post_act=post_act)

    def forward(self, x, *cond_inputs, do_checkpoint=False, **kw_cond_inputs):
        if self._native and not do_checkpoint:
            return self._forward_native(x, cond_inputs, kw_cond_inputs)
        if do_checkpoint:
            dx = checkpoint(self.conv_blocks, x, *cond_inputs, **kw_cond_inputs)
        else:
            dx = self.conv_blocks(x, *cond_inputs, **kw_cond_inputs)
        x_shortcut = self.conv_block_s(x, *cond_inputs, **kw_cond_inputs) if self.learn_shortcut else x
        return x_shortcut + dx


class Res2dBlock(_BaseResBlock):
    """ref: imaginaire/layers/residual.py:305-382"""

    def __init__(self, in_channels, out_channels, kernel_size=3,
                 padding=1, dilation=1, groups=1, bias=True,
                 padding_mode='zeros',
                 weight_norm_type='none', weight_norm_params=None,
                 activation_norm_type='none', activation_norm_params=None,
                 skip_activation_norm=True, skip_nonlinearity=False,
                 nonlinearity='leakyrelu', inplace_nonlinearity=False,
                 apply_noise=False, hidden_channels_equal_out_channels=False,
                 order='CNACNA', learn_shortcut=False):
        super().__init__(in_channels, out_channels, kernel_size, padding,
                         dilation, groups, bias, padding_mode,
                         weight_norm_type, weight_norm_params,
                         activation_norm_type, activation_norm_params,
                         skip_activation_norm, skip_nonlinearity,
                         nonlinearity, inplace_nonlinearity, apply_noise,
                         hidden_channels_equal_out_channels,
                         order, Conv2dBlock, learn_shortcut)
