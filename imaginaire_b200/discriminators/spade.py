"""SPADE discriminator: FPSE + multi-resolution patch discriminators.
ref: imaginaire/discriminators/spade.py:15-117

``Discriminator(dis_cfg, data_cfg).forward(data, net_G_output)`` returns the reference's
dict (real/fake outputs and feature lists).  Every convolution goes through the native
Conv2dBlock.
"""
import torch
from torch import nn

from .. import functional as NF
from ..layers.conv import prefetch_spectral_norms
from ..utils.data import get_paired_input_image_channel_number, get_paired_input_label_channel_number
from .fpse import FPSEDiscriminator
from .multires_patch import NLayerPatchDiscriminator


class Discriminator(nn.Module):
    def __init__(self, dis_cfg, data_cfg):
        super().__init__()
        image_channels = get_paired_input_image_channel_number(data_cfg)
        video = getattr(data_cfg, 'type', '') == 'imaginaire.datasets.paired_videos'
        num_labels = get_paired_input_label_channel_number(data_cfg, video=video)
        kernel_size = getattr(dis_cfg, 'kernel_size', 3)
        num_filters = getattr(dis_cfg, 'num_filters', 128)
        max_num_filters = getattr(dis_cfg, 'max_num_filters', 512)
        num_discriminators = getattr(dis_cfg, 'num_discriminators', 2)
        num_layers = getattr(dis_cfg, 'num_layers', 5)
        activation_norm_type = getattr(dis_cfg, 'activation_norm_type', 'none')
        weight_norm_type = getattr(dis_cfg, 'weight_norm_type', 'spectral')
        self.discriminators = nn.ModuleList([
            NLayerPatchDiscriminator(kernel_size, image_channels + num_labels, num_filters, num_layers,
                                     max_num_filters, activation_norm_type, weight_norm_type)
            for _ in range(num_discriminators)])
        self.fpse_discriminator = FPSEDiscriminator(
            image_channels, num_labels, num_filters, getattr(dis_cfg, 'fpse_kernel_size', 3),
            weight_norm_type, getattr(dis_cfg, 'fpse_activation_norm_type', 'none'))

    def _single_forward(self, input_label, input_image):
        prefetch_spectral_norms(self)   # every layer below runs exactly once per _single_forward
        input_x = torch.cat((input_label, input_image), 1)
        outputs = list(self.fpse_discriminator(input_image, input_label))
        features = []
        for i, net in enumerate(self.discriminators):
            out, feats = net(input_x)
            outputs.append(out)
            features.append(feats)
            if i + 1 < len(self.discriminators):   # the reference also resizes after the last one; nothing reads it
                input_x = NF.bilinear_half_align_corners(input_x)   # interpolate(scale_factor=0.5, bilinear, align_corners=True)
        return outputs, features

    def forward(self, data, net_G_output):
        out = {}
        out['real_outputs'], out['real_features'] = self._single_forward(data['label'], data['images'])
        out['fake_outputs'], out['fake_features'] = self._single_forward(data['label'],
                                                                         net_G_output['fake_images'])
        return out
