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


class Generator(BaseNetwork):
    """vid2vid generator.  ref: imaginaire/generators/vid2vid.py:39-388

    Same constructor (gen_cfg, data_cfg), attribute / state-dict names (``label_embedding``,
    ``up_i``, ``conv_img``, ``fc``, ``down_first``, ``down_i``, ``res_i``, ``flow_network_temp``,
    ``img_prev_embedding``) and ``forward(data) -> dict`` as the reference.  ``flow_network_temp``
    is constructed as upstream does (the fork has that line commented out, SURVEY.md Appendix C,
    which leaves its temporal path without a flow network).  Pose-data label filtering and
    partial-convolution SPADE are not part of this build.

    All feature maps are NHWC bf16 internally; the returned images / flow / mask are fp32 NCHW.
    """

    def __init__(self, gen_cfg, data_cfg):
        super().__init__()
        from .fs_vid2vid import LabelEmbedder
        from ..layers import LinearBlock
        self.gen_cfg = gen_cfg
        self.data_cfg = data_cfg
        self.num_frames_G = data_cfg.num_frames_G
        self.num_layers = num_layers = getattr(gen_cfg, 'num_layers', 7)
        self.num_downsamples_img = getattr(gen_cfg, 'num_downsamples_img', 4)
        self.num_filters = num_filters = getattr(gen_cfg, 'num_filters', 32)
        self.max_num_filters = getattr(gen_cfg, 'max_num_filters', 1024)
        self.kernel_size = kernel_size = getattr(gen_cfg, 'kernel_size', 3)
        padding = kernel_size // 2
        if hasattr(data_cfg, 'for_pose_dataset'):
            raise NotImplementedError('pose datasets are outside this build')
        self.is_pose_data = False

        self.num_input_channels = num_input_channels = get_paired_input_label_channel_number(data_cfg)
        num_img_channels = get_paired_input_image_channel_number(data_cfg)
        aug_cfg = data_cfg.val.augmentations
        if hasattr(aug_cfg, 'center_crop_h_w'):
            crop_h_w = aug_cfg.center_crop_h_w
        elif hasattr(aug_cfg, 'resize_h_w'):
            crop_h_w = aug_cfg.resize_h_w
        else:
            raise ValueError('Need to specify output size.')
        crop_h, crop_w = (int(v) for v in str(crop_h_w).split(','))
        self.sh = crop_h // (2 ** num_layers)
        self.sw = crop_w // (2 ** num_layers)

        self.z_dim = getattr(gen_cfg, 'style_dims', 256)
        self.use_segmap_as_input = getattr(gen_cfg, 'use_segmap_as_input', False)

        self.emb_cfg = emb_cfg = getattr(gen_cfg, 'embed', None)
        self.use_embed = getattr(emb_cfg, 'use_embed', 'True')
        self.num_downsamples_embed = getattr(emb_cfg, 'num_downsamples', 5)
        if self.use_embed:
            self.label_embedding = LabelEmbedder(emb_cfg, num_input_channels)

        self.flow_cfg = flow_cfg = gen_cfg.flow
        self.spade_combine = getattr(flow_cfg, 'multi_spade_combine', True)
        self.num_multi_spade_layers = getattr(flow_cfg.multi_spade_combine, 'num_layers', 3)
        self.temporal_initialized = False
        self.generate_raw_output = False

        weight_norm_type = getattr(gen_cfg, 'weight_norm_type', 'spectral')
        activation_norm_type = gen_cfg.activation_norm_type
        activation_norm_params = gen_cfg.activation_norm_params
        if self.use_embed and not hasattr(activation_norm_params, 'num_filters'):
            activation_norm_params.num_filters = 0
        nonlinearity = 'leakyrelu'
        self.base_res_block = base_res_block = partial(
            Res2dBlock, kernel_size=kernel_size, padding=padding, weight_norm_type=weight_norm_type,
            activation_norm_type=activation_norm_type, activation_norm_params=activation_norm_params,
            nonlinearity=nonlinearity, order='NACNAC')

        for i in range(num_layers, -1, -1):
            activation_norm_params.cond_dims = self.get_cond_dims(i)
            activation_norm_params.partial = False
            setattr(self, 'up_%d' % i, base_res_block(self.get_num_filters(i + 1), self.get_num_filters(i)))

        self.conv_img = Conv2dBlock(num_filters, num_img_channels, kernel_size, padding=padding,
                                    nonlinearity=nonlinearity, order='AC')
        num_filters = min(self.max_num_filters, num_filters * (2 ** (self.num_layers + 1)))
        if self.use_segmap_as_input:
            self.fc = Conv2dBlock(num_input_channels, num_filters, kernel_size=3, padding=1)
        else:
            self.fc = LinearBlock(self.z_dim, num_filters * self.sh * self.sw)
        self.downsample = nn.AvgPool2d(kernel_size=3, stride=2, padding=1)
        self.init_temporal_network()

    # ------------------------------------------------------------------ construction helpers
    def init_temporal_network(self, cfg_init=None):
        """ref: vid2vid.py:300-352"""
        import numpy as np
        from .fs_vid2vid import LabelEmbedder
        from ..utils.init_weight import weights_init
        num_downsamples_img = self.num_downsamples_img
        self.num_res_blocks = int(np.ceil((self.num_layers - num_downsamples_img) / 2.0) * 2)
        num_img_channels = get_paired_input_image_channel_number(self.data_cfg)
        self.down_first = Conv2dBlock(num_img_channels, self.num_filters, self.kernel_size, padding=self.kernel_size // 2)
        init = (lambda m: m.apply(weights_init(cfg_init.type, cfg_init.gain))) if cfg_init is not None else (lambda m: m)
        init(self.down_first)
        activation_norm_params = self.gen_cfg.activation_norm_params
        for i in range(num_downsamples_img + 1):
            activation_norm_params.cond_dims = self.get_cond_dims(i)
            setattr(self, 'down_%d' % i, init(self.base_res_block(self.get_num_filters(i), self.get_num_filters(i + 1))))
        res_ch = self.get_num_filters(num_downsamples_img + 1)
        activation_norm_params.cond_dims = self.get_cond_dims(num_downsamples_img + 1)
        for i in range(self.num_res_blocks):
            setattr(self, 'res_%d' % i, init(self.base_res_block(res_ch, res_ch)))
        flow_cfg = self.flow_cfg
        self.temporal_initialized = True
        self.generate_raw_output = getattr(flow_cfg, 'generate_raw_output', False) and self.spade_combine
        self.flow_network_temp = init(FlowGenerator(flow_cfg, self.data_cfg))      # upstream line, see class docstring
        self.spade_combine = getattr(flow_cfg, 'multi_spade_combine', True)
        if self.spade_combine:
            emb_cfg = flow_cfg.multi_spade_combine.embed
            self.img_prev_embedding = init(LabelEmbedder(emb_cfg, num_img_channels + 1))

    def get_cond_dims(self, num_downs=0):
        """ref: vid2vid.py:354-370"""
        if not self.use_embed:
            return [self.num_input_channels]
        num_filters = getattr(self.emb_cfg, 'num_filters', 32)
        num_downs = min(num_downs, self.num_downsamples_embed)
        ch = [min(self.max_num_filters, num_filters * (2 ** num_downs))]
        if num_downs < self.num_multi_spade_layers:
            ch = ch * 2
        return ch

    def get_cond_maps(self, label, embedder):
        """ref: vid2vid.py:372-388"""
        if not self.use_embed:
            return [[label]] * (self.num_layers + 1)
        embedded_label = embedder(label)
        return [[m] for m in embedded_label]

    # ------------------------------------------------------------------ forward
    def _up(self, x):
        return NF.upsample2x(x)

    def one_up_conv_layer(self, x, encoded_label, i):
        x = getattr(self, 'up_' + str(i))._forward_native(x, tuple(encoded_label), {})
        if i != 0:
            x = self._up(x)
        return x

    def forward(self, data):
        label = data['label']
        label_prev, img_prev = data['prev_labels'], data['prev_images']
        is_first_frame = img_prev is None
        z = getattr(data, 'z', None)          # as in the reference: a dict has no attribute z -> zeros
        bs, _, h, w = label.size()

        cond_maps_now = self.get_cond_maps(label, self.label_embedding)
        if is_first_frame:
            if self.use_segmap_as_input:
                x_img = torch.nn.functional.interpolate(label, size=(self.sh, self.sw))
                x_img = self.fc.forward_native(NF.to_nhwc(x_img))
            else:
                if z is None:
                    z = torch.zeros(bs, self.z_dim, dtype=label.dtype, device=label.device)
                x_img = NF.to_nhwc(self.fc(z).view(bs, -1, self.sh, self.sw))
            for i in range(self.num_layers, self.num_downsamples_img, -1):
                j = min(self.num_downsamples_embed, i)
                x_img = getattr(self, 'up_' + str(i))._forward_native(x_img, tuple(cond_maps_now[j]), {})
                x_img = self._up(x_img)
        else:
            x_img = self.down_first.forward_native(NF.to_nhwc(img_prev[:, -1]))
            cond_maps_prev = self.get_cond_maps(label_prev[:, -1], self.label_embedding)
            for i in range(self.num_downsamples_img + 1):
                j = min(self.num_downsamples_embed, i)
                x_img = getattr(self, 'down_' + str(i))._forward_native(x_img, tuple(cond_maps_prev[j]), {})
                if i != self.num_downsamples_img:
                    x_img = self.downsample(x_img)
            j = min(self.num_downsamples_embed, self.num_downsamples_img + 1)
            for i in range(self.num_res_blocks):
                cond_maps = cond_maps_prev[j] if i < self.num_res_blocks // 2 else cond_maps_now[j]
                x_img = getattr(self, 'res_' + str(i))._forward_native(x_img, tuple(cond_maps), {})

        flow = mask = img_warp = None
        warp_prev = self.temporal_initialized and not is_first_frame and label_prev.shape[1] == self.num_frames_G - 1
        if warp_prev:
            label_concat = torch.cat([label_prev.reshape(bs, -1, h, w), label], dim=1)
            img_prev_concat = img_prev.reshape(bs, -1, h, w)
            flow, mask = self.flow_network_temp(label_concat, img_prev_concat)
            img_warp = warp_prev_frame(img_prev[:, -1], flow)
            if self.spade_combine:
                img_embed = torch.cat([img_warp, mask], dim=1)
                cond_maps_img = self.get_cond_maps(img_embed, self.img_prev_embedding)
                x_raw_img = None

        for i in range(self.num_downsamples_img, -1, -1):
            j = min(i, self.num_downsamples_embed)
            cond_maps = list(cond_maps_now[j])
            if self.generate_raw_output:
                if i >= self.num_multi_spade_layers - 1:
                    x_raw_img = x_img
                if i < self.num_multi_spade_layers:
                    x_raw_img = self.one_up_conv_layer(x_raw_img, cond_maps, i)
            if warp_prev and i < self.num_multi_spade_layers:
                cond_maps = cond_maps + list(cond_maps_img[j])
            x_img = self.one_up_conv_layer(x_img, cond_maps, i)

        nc = get_paired_input_image_channel_number(self.data_cfg)
        img_final = NF.to_nchw(self.conv_img.forward_native(x_img), nc, apply_tanh=True)
        img_raw = None
        if self.spade_combine and self.generate_raw_output:
            img_raw = NF.to_nchw(self.conv_img.forward_native(x_raw_img), nc, apply_tanh=True)
        if warp_prev and not self.spade_combine:
            img_raw = img_final
            img_final = img_final * mask + img_warp * (1 - mask)
        return {'fake_images': img_final, 'fake_flow_maps': flow, 'fake_occlusion_masks': mask,
                'fake_raw_images': img_raw, 'warped_images': img_warp}
