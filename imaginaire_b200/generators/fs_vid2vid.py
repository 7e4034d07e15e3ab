"""Label / image embedding network of the vid2vid generators on the native kernels.

ref: imaginaire/generators/fs_vid2vid.py:1072-1176 (LabelEmbedder)

The reference builds it from HyperConv2dBlock with ``is_hyper_conv`` False everywhere the vid2vid
generator uses it (num_hyper_layers = 0), i.e. plain convolution blocks; the state-dict keys
(``conv_first.layers.conv.*``, ``down_i...``, ``up_i...``) are those of Conv2dBlock.  Hyper
(per-sample predicted) filters belong to few-shot vid2vid and are refused here.
"""
from functools import partial

import torch
from torch import nn

from .. import functional as NF
from ..layers import Conv2dBlock


class LabelEmbedder(nn.Module):
    def __init__(self, emb_cfg, num_input_channels, num_hyper_layers=0):
        super().__init__()
        if num_hyper_layers != 0:
            raise NotImplementedError('hyper-convolution embedders (few-shot vid2vid) are outside the hot path')
        num_filters = getattr(emb_cfg, 'num_filters', 32)
        max_num_filters = getattr(emb_cfg, 'max_num_filters', 1024)
        self.arch = getattr(emb_cfg, 'arch', 'encoderdecoder')
        self.num_downsamples = num_downsamples = getattr(emb_cfg, 'num_downsamples', 5)
        kernel_size = getattr(emb_cfg, 'kernel_size', 3)
        weight_norm_type = getattr(emb_cfg, 'weight_norm_type', 'spectral')
        activation_norm_type = getattr(emb_cfg, 'activation_norm_type', 'none')
        self.unet = 'unet' in self.arch
        self.has_decoder = 'decoder' in self.arch or self.unet
        self.num_hyper_layers = 0

        base_conv_block = partial(Conv2dBlock, kernel_size=kernel_size, padding=(kernel_size // 2),
                                  weight_norm_type=weight_norm_type, activation_norm_type=activation_norm_type,
                                  nonlinearity='leakyrelu')
        ch = [min(max_num_filters, num_filters * (2 ** i)) for i in range(num_downsamples + 1)]
        self.conv_first = base_conv_block(num_input_channels, num_filters, activation_norm_type='none')
        for i in range(num_downsamples):
            setattr(self, 'down_%d' % i, base_conv_block(ch[i], ch[i + 1], stride=2))
        if self.has_decoder:
            self.upsample = nn.Upsample(scale_factor=2)
            for i in reversed(range(num_downsamples)):
                ch_i = ch[i + 1] * (2 if self.unet and i != num_downsamples - 1 else 1)
                setattr(self, 'up_%d' % i, base_conv_block(ch_i, ch[i]))

    def forward(self, input, weights=None):
        """Returns the list of feature maps (finest first), NHWC bf16 internal tensors: they are
        consumed as SPADE conditioning inputs by the native residual blocks."""
        if input is None:
            return None
        if weights is not None:
            raise NotImplementedError('hyper-convolution weights are not supported')
        output = [self.conv_first.forward_native(NF.to_nhwc(input))]
        for i in range(self.num_downsamples):
            output.append(getattr(self, 'down_%d' % i).forward_native(output[-1]))
        if not self.has_decoder:
            return output
        if not self.unet:
            output = [output[-1]]
        for i in reversed(range(self.num_downsamples)):
            input_i = output[-1]
            if self.unet and i != self.num_downsamples - 1:
                input_i = torch.cat([input_i, output[i + 1]], dim=1)
            input_i = NF.upsample2x(input_i)
            output.append(getattr(self, 'up_%d' % i).forward_native(input_i))
        if self.unet:
            output = output[self.num_downsamples:]
        return output[::-1]
