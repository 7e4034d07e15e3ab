"""The flow branch of the vid2vid generator on the native kernels.

ref: imaginaire/generators/vid2vid.py:390-481 (FlowGenerator), :27-36 (get_num_filters),
     :222-230 (flow estimation + warp of the previous frame inside Generator.forward)

``FlowGenerator`` keeps the reference's constructor, attribute names and state-dict keys
(``down_lbl.N.layers.*``, ``down_img.N...``, ``res_flow.N.conv_block_K...``, ``up_flow.(2k+1)...``,
``conv_flow.0...``, ``conv_mask.0...``), so a reference checkpoint's ``flow_network_temp.*``
entries load with ``strict=True``.  Every convolution / normalisation runs through the tcgen05
implicit-GEMM and fused norm kernels in NHWC bf16; ``warp_prev_frame`` is the generator's
``resample(img_prev[:, -1], flow)`` on the bit-exact grid_sample-form warp kernel.

The embedding networks and the SPADE up-blocks of the full vid2vid ``Generator`` are built from
the same ``Conv2dBlock`` / ``Res2dBlock`` / ``SpatiallyAdaptiveNorm`` classes (multi-cond lists,
``num_filters: 0`` 1x1 projections: tests/test_gpu_variants.py); assembling that generator and
its trainer is the next row.
"""
from functools import partial

import torch
from torch import nn

from .. import functional as NF
from ..layers import Conv2dBlock, Res2dBlock
from ..model_utils.fs_vid2vid import resample
from ..utils.data import get_paired_input_image_channel_number, get_paired_input_label_channel_number


class BaseNetwork(nn.Module):
    """ref: imaginaire/generators/vid2vid.py:20-36"""

    def get_num_filters(self, num_downsamples):
        return min(self.max_num_filters, self.num_filters * (2 ** num_downsamples))


class FlowGenerator(BaseNetwork):
    """Estimates the flow from the previous frame(s) and the occlusion mask.

    flow_cfg: ``num_filters`` (32), ``max_num_filters`` (1024), ``num_downsamples`` (5),
    ``kernel_size`` (3), ``num_res_blocks`` (6), ``flow_output_multiplier`` (20),
    ``activation_norm_type`` ('sync_batch'), ``weight_norm_type`` ('spectral').
    """

    def __init__(self, flow_cfg, data_cfg):
        super().__init__()
        num_input_channels = get_paired_input_label_channel_number(data_cfg)
        num_prev_img_channels = get_paired_input_image_channel_number(data_cfg)
        num_frames = data_cfg.num_frames_G

        self.num_filters = num_filters = getattr(flow_cfg, 'num_filters', 32)
        self.max_num_filters = getattr(flow_cfg, 'max_num_filters', 1024)
        num_downsamples = getattr(flow_cfg, 'num_downsamples', 5)
        kernel_size = getattr(flow_cfg, 'kernel_size', 3)
        padding = kernel_size // 2
        self.num_res_blocks = getattr(flow_cfg, 'num_res_blocks', 6)
        self.flow_output_multiplier = getattr(flow_cfg, 'flow_output_multiplier', 20)
        activation_norm_type = getattr(flow_cfg, 'activation_norm_type', 'sync_batch')
        weight_norm_type = getattr(flow_cfg, 'weight_norm_type', 'spectral')

        base_conv_block = partial(Conv2dBlock, kernel_size=kernel_size, padding=padding,
                                  weight_norm_type=weight_norm_type,
                                  activation_norm_type=activation_norm_type, nonlinearity='leakyrelu')

        # Same construction order as the reference, so a seeded build draws the same weights.
        down_lbl = [base_conv_block(num_input_channels * num_frames, num_filters)]
        down_img = [base_conv_block(num_prev_img_channels * (num_frames - 1), num_filters)]
        for i in range(num_downsamples):
            down_lbl += [base_conv_block(self.get_num_filters(i), self.get_num_filters(i + 1), stride=2)]
            down_img += [base_conv_block(self.get_num_filters(i), self.get_num_filters(i + 1), stride=2)]

        res_flow = []
        ch = self.get_num_filters(num_downsamples)
        for i in range(self.num_res_blocks):
            res_flow += [Res2dBlock(ch, ch, kernel_size, padding=padding, weight_norm_type=weight_norm_type,
                                    activation_norm_type=activation_norm_type, order='CNACN')]

        up_flow = []
        for i in reversed(range(num_downsamples)):
            up_flow += [nn.Upsample(scale_factor=2),
                        base_conv_block(self.get_num_filters(i + 1), self.get_num_filters(i))]

        conv_flow = [Conv2dBlock(num_filters, 2, kernel_size, padding=padding)]
        conv_mask = [Conv2dBlock(num_filters, 1, kernel_size, padding=padding, nonlinearity='sigmoid')]

        self.down_lbl = nn.Sequential(*down_lbl)
        self.down_img = nn.Sequential(*down_img)
        self.res_flow = nn.Sequential(*res_flow)
        self.up_flow = nn.Sequential(*up_flow)
        self.conv_flow = nn.Sequential(*conv_flow)
        self.conv_mask = nn.Sequential(*conv_mask)

    @staticmethod
    def _run(seq, x):
        for m in seq:
            if isinstance(m, nn.Upsample):
                x = NF.upsample2x(x)               # nearest x2 in NHWC bf16 (the reference's default mode)
            elif isinstance(m, Res2dBlock):
                x = m._forward_native(x, (), {})
            else:
                x = m.forward_native(x)
        return x

    def forward(self, label, img_prev):
        """label: B x (L * num_frames) x H x W, img_prev: B x (3 * (num_frames - 1)) x H x W.
        Returns (flow B x 2 x H x W, mask B x 1 x H x W), fp32 NCHW."""
        downsample = self._run(self.down_lbl, NF.to_nhwc(label)) + self._run(self.down_img, NF.to_nhwc(img_prev))
        res = self._run(self.res_flow, downsample)
        flow_feat = self._run(self.up_flow, res)
        flow = NF.to_nchw(self._run(self.conv_flow, flow_feat), 2) * self.flow_output_multiplier
        mask = NF.to_nchw(self._run(self.conv_mask, flow_feat), 1)
        return flow, mask


def warp_prev_frame(img_prev_last, flow):
    """ref: imaginaire/generators/vid2vid.py:230 ``img_warp = resample(img_prev[:, -1], flow)``"""
    return resample(img_prev_last.float().contiguous(), flow.float().contiguous())
