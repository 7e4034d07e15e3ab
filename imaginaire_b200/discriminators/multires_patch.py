"""N-layer patch discriminator.  ref: imaginaire/discriminators/multires_patch.py:250-313

Built from the native Conv2dBlock (k=4 stride-2 / stride-1 tcgen05 convolutions with the
LeakyReLU in the epilogue); same layer names (layer0..layerN) and state-dict keys.
"""
import functools

import numpy as np
from torch import nn

from ..layers import Conv2dBlock


class NLayerPatchDiscriminator(nn.Module):
    def __init__(self, kernel_size, num_input_channels, num_filters, num_layers, max_num_filters,
                 activation_norm_type, weight_norm_type):
        super().__init__()
        self.num_layers = num_layers
        padding = int(np.floor((kernel_size - 1.0) / 2))
        block = functools.partial(Conv2dBlock, kernel_size=kernel_size, padding=padding,
                                  weight_norm_type=weight_norm_type,
                                  activation_norm_type=activation_norm_type,
                                  nonlinearity='leakyrelu', order='CNA')
        layers = [block(num_input_channels, num_filters, stride=2)]
        for n in range(num_layers):
            prev, num_filters = num_filters, min(num_filters * 2, max_num_filters)
            layers.append(block(prev, num_filters, stride=2 if n < num_layers - 1 else 1))
        layers.append(Conv2dBlock(num_filters, 1, 3, 1, padding, weight_norm_type=weight_norm_type))
        for n, layer in enumerate(layers):
            setattr(self, 'layer%d' % n, nn.Sequential(layer))

    def forward(self, input_x):
        res = [input_x]
        for n in range(self.num_layers + 2):
            res.append(getattr(self, 'layer%d' % n)(res[-1]))
        return res[-1], res[1:-1]
