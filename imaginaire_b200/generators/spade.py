"""SPADE / GauGAN generator with the reference's API, built from the native blocks.

ref: imaginaire/generators/spade.py:22-214 (Generator), :217-493 (SPADEGenerator),
     :496-563 (StyleEncoder)

``Generator(gen_cfg, data_cfg)``; ``forward(data, random_style=False)`` returns
``{'fake_images', 'mu', 'logvar'}``; ``inference(...)`` returns ``(images, file_names)``.
Sub-module names, construction order and parameter names are the reference's, so a
seeded build (plus the trainer's ``weights_init``) reproduces the reference's weights bit
for bit and its checkpoints load with ``strict=True``.

Inside, activations are NHWC bf16 from the head convolution to the to-RGB convolution;
``fake_images`` is returned as NCHW fp32 like the reference's.  Differences in
execution only: the nearest x2 upsample, normalisation statistics, modulation,
activations and residual adds are fused into neighbouring kernels (see layers/), the
label map is resized once per resolution instead of once per SPADE norm, and the
positional-encoding grid is a non-persistent buffer instead of a ``.cuda()`` tensor.
"""
import functools
import math
import types

import numpy as np
import torch
import torch.nn.functional as F
from torch import nn

from .. import functional as NF
from ..layers import Conv2dBlock, LinearBlock, Res2dBlock
from ..layers.conv import prefetch_spectral_norms
from ..layers.residual import _inner_stats_mode
from ..utils.data import (get_crop_h_w, get_paired_input_image_channel_number,
                          get_paired_input_label_channel_number)


class Generator(nn.Module):
    def __init__(self, gen_cfg, data_cfg):
        super().__init__()
        image_channels = get_paired_input_image_channel_number(data_cfg)
        num_labels = get_paired_input_label_channel_number(data_cfg)
        crop_h, crop_w = get_crop_h_w(data_cfg.train.augmentations)
        small_side = min(crop_w, crop_h)
        num_filters = getattr(gen_cfg, 'num_filters', 128)
        kernel_size = getattr(gen_cfg, 'kernel_size', 3)
        weight_norm_type = getattr(gen_cfg, 'weight_norm_type', 'spectral')

        cond_dims = 0
        self.style_dims = getattr(gen_cfg, 'style_dims', None)
        self.use_style = self.style_dims is not None
        if self.use_style:
            cond_dims += self.style_dims
        self.use_attribute = hasattr(gen_cfg, 'attribute_dims')
        if self.use_attribute:
            self.attribute_dims = gen_cfg.attribute_dims
            cond_dims += gen_cfg.attribute_dims
        self.use_style_encoder = self.use_style or self.use_attribute

        skip_activation_norm = getattr(gen_cfg, 'skip_activation_norm', True)
        an_params = getattr(gen_cfg, 'activation_norm_params', None)
        if an_params is None:
            an_params = types.SimpleNamespace()
        # defaults of the SPADE norm (spade.py:74-93)
        for key, default in (('num_filters', 128), ('kernel_size', 3),
                             ('activation_norm_type', 'sync_batch'), ('separate_projection', False)):
            if not hasattr(an_params, key):
                setattr(an_params, key, default)
        if not hasattr(an_params, 'activation_norm_params'):
            an_params.activation_norm_params = types.SimpleNamespace(affine=True)
        setattr(an_params, 'cond_dims', num_labels)
        if not hasattr(an_params, 'weight_norm_type'):
            setattr(an_params, 'weight_norm_type', weight_norm_type)
        global_adaptive_norm_type = getattr(gen_cfg, 'global_adaptive_norm_type', 'sync_batch')
        use_posenc = getattr(gen_cfg, 'use_posenc_in_input_layer', True)

        self.spade_generator = SPADEGenerator(num_labels, small_side, image_channels, num_filters,
                                              kernel_size, cond_dims, an_params, weight_norm_type,
                                              global_adaptive_norm_type, skip_activation_norm,
                                              use_posenc, self.use_style_encoder)
        if self.use_style:
            enc_cfg = getattr(gen_cfg, 'style_enc', None)
            if enc_cfg is None:
                enc_cfg = types.SimpleNamespace()
            for key, default in (('num_filters', 128), ('kernel_size', 3), ('freeze_random', False),
                                 ('weight_norm_type', weight_norm_type)):
                if not hasattr(enc_cfg, key):
                    setattr(enc_cfg, key, default)
            setattr(enc_cfg, 'input_image_channels', image_channels)
            setattr(enc_cfg, 'style_dims', self.style_dims)
            self.style_encoder = StyleEncoder(enc_cfg)
        self.z = None

    def _random_z(self, label):
        return torch.randn(label.size(0), self.style_dims, dtype=torch.float32, device=label.device)

    def forward(self, data, random_style=False):
        prefetch_spectral_norms(self.spade_generator)
        if self.use_style_encoder:
            if not random_style:
                prefetch_spectral_norms(self.style_encoder)
            if random_style:
                z, mu, logvar = self._random_z(data['label']), None, None
            else:
                mu, logvar, z = self.style_encoder(data['images'])
            data['z'] = torch.cat((z, data['attributes'].squeeze(1)), dim=1) if self.use_attribute else z
        output = self.spade_generator(data)
        if self.use_style_encoder:
            output['mu'] = mu
            output['logvar'] = logvar
        return output

    def inference(self, data, random_style=False, use_fixed_random_style=False, keep_original_size=False):
        self.eval()
        self.spade_generator.eval()
        if random_style:
            if self.z is None or not use_fixed_random_style:
                self.z = self._random_z(data['label'])
            z = self.z
        else:
            _, _, z = self.style_encoder(data['images'])
        data['z'] = z
        images = self.spade_generator(data)['fake_images']
        if keep_original_size:
            height, width = data['original_h_w'][0][0], data['original_h_w'][0][1]
            images = F.interpolate(images, size=[height, width])
        return images, data['key']['seg_maps'][0]


# Debugging aid (tools/generator_error_trace.py): set to a list to receive (block name, NHWC output) after every
# block of SPADEGenerator.forward.
TRACE = None


class SPADEGenerator(nn.Module):
    def __init__(self, num_labels, out_image_small_side_size, image_channels, num_filters,
                 kernel_size, style_dims, activation_norm_params, weight_norm_type,
                 global_adaptive_norm_type, skip_activation_norm, use_posenc_in_input_layer,
                 use_style_encoder):
        super().__init__()
        self.use_style_encoder = use_style_encoder
        self.use_posenc_in_input_layer = use_posenc_in_input_layer
        self.out_image_small_side_size = out_image_small_side_size
        self.num_filters = num_filters
        self.image_channels = image_channels
        padding = int(np.ceil((kernel_size - 1.0) / 2))
        act = 'leakyrelu'
        res = functools.partial(Res2dBlock, kernel_size=kernel_size, padding=padding,
                                bias=[True, True, False], weight_norm_type=weight_norm_type,
                                activation_norm_type='spatially_adaptive',
                                activation_norm_params=activation_norm_params,
                                skip_activation_norm=skip_activation_norm,
                                nonlinearity=act, order='NACNAC')
        if use_style_encoder:
            self.fc_0 = LinearBlock(style_dims, 2 * style_dims, weight_norm_type=weight_norm_type,
                                    nonlinearity='relu', order='CAN')
            self.fc_1 = LinearBlock(2 * style_dims, 2 * style_dims, weight_norm_type=weight_norm_type,
                                    nonlinearity='relu', order='CAN')
            adaptive = types.SimpleNamespace(
                cond_dims=2 * style_dims, activation_norm_type=global_adaptive_norm_type,
                weight_norm_type=activation_norm_params.weight_norm_type,
                separate_projection=activation_norm_params.separate_projection)
            adaptive.activation_norm_params = types.SimpleNamespace(
                affine=activation_norm_params.activation_norm_params.affine)
            mid = functools.partial(Conv2dBlock, kernel_size=kernel_size, stride=1, padding=padding,
                                    bias=True, weight_norm_type=weight_norm_type,
                                    activation_norm_type='adaptive', activation_norm_params=adaptive,
                                    nonlinearity=act, order='NAC')
            mid_prefix = 'cbn_'
        else:
            mid = functools.partial(Conv2dBlock, kernel_size=kernel_size, stride=1, padding=padding,
                                    bias=True, weight_norm_type=weight_norm_type,
                                    nonlinearity=act, order='NAC')
            mid_prefix = 'conv_'
        self._mid_prefix = mid_prefix
        f = num_filters
        in_labels = num_labels + (2 if use_posenc_in_input_layer else 0)
        self.head_0 = Conv2dBlock(in_labels, 8 * f, kernel_size=kernel_size, stride=1, padding=padding,
                                  weight_norm_type=weight_norm_type, activation_norm_type='none',
                                  nonlinearity=act)
        setattr(self, mid_prefix + 'head_0', mid(8 * f, 16 * f))
        self.head_1 = res(16 * f, 16 * f)
        self.head_2 = res(16 * f, 16 * f)
        self.up_0a = res(16 * f, 8 * f)
        setattr(self, mid_prefix + 'up_0a', mid(8 * f, 8 * f))
        self.up_0b = res(8 * f, 8 * f)
        self.up_1a = res(8 * f, 4 * f)
        setattr(self, mid_prefix + 'up_1a', mid(4 * f, 4 * f))
        self.up_1b = res(4 * f, 4 * f)
        self.up_2a = res(4 * f, 4 * f)
        setattr(self, mid_prefix + 'up_2a', mid(4 * f, 4 * f))
        self.up_2b = res(4 * f, 2 * f)
        to_rgb = functools.partial(Conv2dBlock, kernel_size=5, stride=1, padding=2,
                                   weight_norm_type=weight_norm_type, activation_norm_type='none',
                                   nonlinearity=act, order='ANC')
        self.conv_img256 = to_rgb(2 * f, image_channels)
        self.base = 16
        if out_image_small_side_size == 512:
            self.up_3a = res(2 * f, f)
            self.up_3b = res(f, f)
            self.conv_img512 = to_rgb(f, image_channels)
            self.base = 32
        if out_image_small_side_size == 1024:
            self.up_3a = res(2 * f, f)
            self.up_3b = res(f, f)
            self.up_4a = res(f, f // 2)
            self.up_4b = res(f // 2, f // 2)
            self.conv_img1024 = to_rgb(f // 2, image_channels)
            self.base = 64
        if out_image_small_side_size not in (256, 512, 1024):
            raise ValueError('Generation image size (%d, %d) not supported' %
                             (out_image_small_side_size, out_image_small_side_size))
        # positional-encoding grid (spade.py:396-399); non-persistent so state-dict keys are unchanged
        xv, yv = torch.meshgrid([torch.arange(-1, 1.1, 2. / 15), torch.arange(-1, 1.1, 2. / 15)], indexing='ij')
        self.register_buffer('xy', torch.cat((xv.unsqueeze(0), yv.unsqueeze(0)), 0).unsqueeze(0),
                             persistent=False)

    def _mid(self, name):
        return getattr(self, self._mid_prefix + name)

    def _run(self, blocks, x, seg, z, final_act=None):
        """Runs a chain of blocks; every block is told which statistics its consumer needs so
        its last convolution can emit them from the epilogue.  final_act: activation of the chain's
        consumer, applied in the epilogue of the last residual block's final convolution."""
        if final_act is not None:
            blocks[-1][0].post_act = final_act
        for k, (blk, kind) in enumerate(blocks):
            nxt = blocks[k + 1][0] if k + 1 < len(blocks) else None
            mode = None
            if nxt is not None:
                first = nxt.conv_block_0 if isinstance(nxt, Res2dBlock) else nxt
                mode = _inner_stats_mode(first)
            if isinstance(blk, Res2dBlock):
                blk.emit_stats = mode
                x = blk(x, seg)
            elif kind == 'style':
                x = blk.forward_native(x, (z,), None, stats_mode=mode)
            else:
                x = blk.forward_native(x, (), None, stats_mode=mode)
            if TRACE is not None:
                TRACE.append((self._block_name(blk), x))
        return x

    def _block_name(self, blk):
        for name, child in self.named_children():
            if child is blk:
                return name
        return type(blk).__name__

    def forward(self, data):
        seg = data['label']
        z = None
        if self.use_style_encoder:
            z = self.fc_1(self.fc_0(data['z']))
        mid_kind = 'style' if self.use_style_encoder else 'plain'

        sy = math.floor(seg.size(2) * 1.0 / self.base)
        sx = math.floor(seg.size(3) * 1.0 / self.base)
        in_seg = F.interpolate(seg, size=[sy, sx], mode='nearest')
        if self.use_posenc_in_input_layer:
            in_xy = F.interpolate(self.xy.to(seg.dtype), size=[sy, sx], mode='bicubic')
            in_seg = torch.cat((in_seg, in_xy.expand(in_seg.size(0), 2, sy, sx)), 1)

        up = NF.upsample2x
        x = self._run([(self.head_0, 'plain'), (self._mid('head_0'), mid_kind), (self.head_1, 'res'),
                       (self.head_2, 'res')], in_seg, seg, z)
        x = up(x)
        x = self._run([(self.up_0a, 'res'), (self._mid('up_0a'), mid_kind), (self.up_0b, 'res')], x, seg, z)
        x = up(x)
        x = self._run([(self.up_1a, 'res'), (self._mid('up_1a'), mid_kind), (self.up_1b, 'res')], x, seg, z)
        x = up(x)
        c = self.image_channels
        # 256-pixel output: the to-RGB block (LeakyReLU -> conv) is the only consumer of up_2b's
        # output, and LeakyReLU commutes with the nearest x2 upsampling in between, so it is applied
        # in the epilogue of up_2b's last convolution (at a quarter of the pixels) instead of as a
        # pass of its own over the 256-channel full-resolution tensor.
        rgb_act = self.conv_img256.leading_act_code() if self.out_image_small_side_size == 256 else None
        if rgb_act is not None and not (isinstance(self.up_2b, Res2dBlock) and self.up_2b._native
                                        and self.up_2b.conv_block_1.ends_with_plain_conv()):
            rgb_act = None
        x = self._run([(self.up_2a, 'res'), (self._mid('up_2a'), mid_kind), (self.up_2b, 'res')], x, seg, z,
                      final_act=rgb_act)
        x = up(x)
        if self.out_image_small_side_size == 256:
            out = NF.to_nchw(self.conv_img256.forward_native(x, skip_first_act=rgb_act is not None), c, apply_tanh=True)
        else:
            x256 = up(self.conv_img256.forward_native(x))
            x = self._run([(self.up_3a, 'res'), (self.up_3b, 'res')], x, seg, z)
            x = up(x)
            x512 = self.conv_img512.forward_native(x)
            if self.out_image_small_side_size == 512:
                out = torch.tanh(NF.to_nchw(x256, c) + NF.to_nchw(x512, c))
            else:
                x256 = up(x256)
                x512 = up(x512)
                x = self._run([(self.up_4a, 'res'), (self.up_4b, 'res')], x, seg, z)
                x = up(x)
                x1024 = self.conv_img1024.forward_native(x)
                out = torch.tanh(NF.to_nchw(x256, c) + NF.to_nchw(x512, c) + NF.to_nchw(x1024, c))
        return {'fake_images': out}


class StyleEncoder(nn.Module):
    """ref: imaginaire/generators/spade.py:496-563"""

    def __init__(self, style_enc_cfg):
        super().__init__()
        cin = style_enc_cfg.input_image_channels
        f = style_enc_cfg.num_filters
        k = style_enc_cfg.kernel_size
        padding = int(np.ceil((k - 1.0) / 2))
        conv = functools.partial(Conv2dBlock, kernel_size=k, stride=2, padding=padding,
                                 weight_norm_type=style_enc_cfg.weight_norm_type,
                                 activation_norm_type='none', nonlinearity='leakyrelu')
        self.layer1 = conv(cin, f)
        self.layer2 = conv(f * 1, f * 2)
        self.layer3 = conv(f * 2, f * 4)
        self.layer4 = conv(f * 4, f * 8)
        self.layer5 = conv(f * 8, f * 8)
        self.layer6 = conv(f * 8, f * 8)
        self.fc_mu = LinearBlock(f * 8 * 4 * 4, style_enc_cfg.style_dims)
        self.fc_var = LinearBlock(f * 8 * 4 * 4, style_enc_cfg.style_dims)
        self.freeze_random = style_enc_cfg.freeze_random
        self.eps = None

    def forward(self, input_x):
        if input_x.size(2) != 256 or input_x.size(3) != 256:
            input_x = F.interpolate(input_x, size=(256, 256), mode='bilinear')
        x = input_x
        for layer in (self.layer1, self.layer2, self.layer3, self.layer4, self.layer5, self.layer6):
            x = layer(x)
        # flatten in the reference's NCHW order (fc weights are indexed c*16 + y*4 + x)
        x = NF.to_nchw(x).reshape(x.size(0), -1)
        mu = self.fc_mu(x)
        logvar = self.fc_var(x)
        std = torch.exp(0.5 * logvar)
        if self.eps is None or not self.freeze_random:
            self.eps = torch.randn_like(std)
        z = self.eps.mul(std) + mu
        return mu, logvar, z
