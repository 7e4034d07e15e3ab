"""Feature-pyramid semantics-embedding discriminator.
ref: imaginaire/discriminators/fpse.py:15-132

Convolutions run on the native kernels; the few glue operations between them (bilinear
x2 upsample, 2x2 average pooling, the label-embedding inner product) are torch
elementwise / pooling calls on the bf16 NHWC tensors.
"""
import functools

import numpy as np
import torch
import torch.nn.functional as F
from torch import nn

from .. import functional as NF
from ..layers import Conv2dBlock


def _pooled_label_maps(segmaps, levels):
    """NHWC bf16 label maps average-pooled by 4, 8, ... (one per pyramid level).  The discriminator
    sees the same label tensor for the real and the fake pass, in the D and in the G update: the
    pyramid is kept on the tensor object (it dies with the batch) and rebuilt if the tensor changes."""
    cache = segmaps.__dict__.get('_imx_fpse_pooled') if not segmaps.requires_grad else None
    if cache is not None and cache[0] == segmaps._version and len(cache[1]) == levels:
        return cache[1]
    pooled = F.avg_pool2d(segmaps.float(), kernel_size=2, stride=2)
    maps = []
    for _ in range(levels):
        pooled = F.avg_pool2d(pooled, kernel_size=2, stride=2)
        maps.append(NF.to_nhwc(pooled))
    if not segmaps.requires_grad:
        segmaps.__dict__['_imx_fpse_pooled'] = (segmaps._version, maps)
    return maps


class FPSEDiscriminator(nn.Module):
    def __init__(self, num_input_channels, num_labels, num_filters, kernel_size, weight_norm_type,
                 activation_norm_type):
        super().__init__()
        padding = int(np.ceil((kernel_size - 1.0) / 2))
        common = dict(weight_norm_type=weight_norm_type, activation_norm_type=activation_norm_type,
                      nonlinearity='leakyrelu', order='CNA')
        same = functools.partial(Conv2dBlock, kernel_size=kernel_size, stride=1, padding=padding, **common)
        down = functools.partial(Conv2dBlock, kernel_size=kernel_size, stride=2, padding=padding, **common)
        lat = functools.partial(Conv2dBlock, kernel_size=1, stride=1, **common)
        f = num_filters
        self.enc1 = down(num_input_channels, f)
        self.enc2 = down(1 * f, 2 * f)
        self.enc3 = down(2 * f, 4 * f)
        self.enc4 = down(4 * f, 8 * f)
        self.enc5 = down(8 * f, 8 * f)
        self.lat2 = lat(2 * f, 4 * f)
        self.lat3 = lat(4 * f, 4 * f)
        self.lat4 = lat(8 * f, 4 * f)
        self.lat5 = lat(8 * f, 4 * f)
        self.upsample2x = nn.Upsample(scale_factor=2, mode='bilinear', align_corners=False)
        self.final2 = same(4 * f, 2 * f)
        self.final3 = same(4 * f, 2 * f)
        self.final4 = same(4 * f, 2 * f)
        self.output = Conv2dBlock(f * 2, 1, kernel_size=1)
        self.seg = Conv2dBlock(f * 2, f * 2, kernel_size=1)
        self.embedding = Conv2dBlock(num_labels, f * 2, kernel_size=1)

    def forward(self, images, segmaps):
        feat11 = self.enc1(images)
        feat12 = self.enc2(feat11)
        feat13 = self.enc3(feat12)
        feat14 = self.enc4(feat13)
        feat15 = self.enc5(feat14)
        feat25 = self.lat5(feat15)
        # upsample2x (bilinear, align_corners=False) + lateral in one kernel; self.upsample2x stays as
        # the reference's attribute
        feat24 = NF.bilinear_up2x_add(feat25, self.lat4(feat14))
        feat23 = NF.bilinear_up2x_add(feat24, self.lat3(feat13))
        feat22 = NF.bilinear_up2x_add(feat23, self.lat2(feat12))
        feats = (self.final2(feat22), self.final3(feat23), self.final4(feat24))
        # The reference embeds the full-resolution label map (1x1 conv to 2F channels) and then
        # average-pools the embedding 4 times.  A 1x1 convolution commutes with average pooling,
        # so the same embeddings are computed from the pooled label maps: a 2F-channel tensor is
        # never materialised at full resolution.  The embedding filter (and its spectral-norm
        # power iteration) is evaluated once per forward, as in the reference.
        conv = self.embedding.layers['conv']
        weight = conv.effective_weight()
        preds = []
        for feat, pooled in zip(feats, _pooled_label_maps(segmaps, len(feats))):
            segemb = NF.conv2d(pooled, weight, conv.bias, None, 1, 0)
            pred = self.output(feat) + NF.channel_dot(segemb, self.seg(feat))
            preds.append(pred)
        return tuple(preds)
