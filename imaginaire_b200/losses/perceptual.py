"""VGG-19 perceptual loss.  ref: imaginaire/losses/perceptual.py:15-191

The feature extractor is VGG-19 up to relu_5_1 built from the native Conv2dBlock (3x3
tcgen05 convolutions with the ReLU in the epilogue, native 2x2 max-pools).  The reference
downloads torchvision's ImageNet weights (perceptual.py:177); here they come from a local
file: ``cfg.trainer.perceptual_loss.weights_path`` or the environment variable
``IMX_VGG19_WEIGHTS`` names a torchvision ``vgg19`` state dict (``features.N.weight/bias`` or
the bare ``N.weight/bias`` of ``vgg19().features``), remapped onto the blocks by
``load_torchvision_state_dict``.  Without one the filters stay randomly initialised and a
warning says so once (fine for throughput runs, not the reference's loss).
"""
import os
import warnings

import torch
import torch.nn.functional as F
from torch import nn

from .. import functional as NF
from ..layers import Conv2dBlock

_VGG19 = [('relu_1_1', 3, 64), ('relu_1_2', 64, 64), 'pool',
          ('relu_2_1', 64, 128), ('relu_2_2', 128, 128), 'pool',
          ('relu_3_1', 128, 256), ('relu_3_2', 256, 256), ('relu_3_3', 256, 256), ('relu_3_4', 256, 256), 'pool',
          ('relu_4_1', 256, 512), ('relu_4_2', 512, 512), ('relu_4_3', 512, 512), ('relu_4_4', 512, 512), 'pool',
          ('relu_5_1', 512, 512)]


class _VGG19Features(nn.Module):
    def __init__(self, layers):
        super().__init__()
        self.wanted = list(layers)
        self.names = []
        self.blocks = nn.ModuleList()
        for item in _VGG19:
            if item == 'pool':
                self.names.append('pool')
                self.blocks.append(nn.Identity())
            else:
                name, cin, cout = item
                self.names.append(name)
                self.blocks.append(Conv2dBlock(cin, cout, 3, padding=1, nonlinearity='relu'))
        last = max(self.names.index(n) for n in self.wanted)
        self.names, self.blocks = self.names[:last + 1], self.blocks[:last + 1]
        for p in self.parameters():
            p.requires_grad = False
        self.pretrained = False

    # torchvision's vgg19().features: index of the Conv2d that precedes each named ReLU (perceptual.py:178-190
    # names the ReLU outputs 1, 3, 6, ...; the convolution sits one index earlier)
    _TV_CONV_INDEX = {'relu_1_1': 0, 'relu_1_2': 2, 'relu_2_1': 5, 'relu_2_2': 7, 'relu_3_1': 10, 'relu_3_2': 12,
                      'relu_3_3': 14, 'relu_3_4': 16, 'relu_4_1': 19, 'relu_4_2': 21, 'relu_4_3': 23,
                      'relu_4_4': 25, 'relu_5_1': 28}

    def load_torchvision_state_dict(self, sd):
        """Copies ``features.N.{weight,bias}`` (or ``N.{weight,bias}``) of a torchvision VGG-19 onto the blocks."""
        prefix = 'features.' if any(k.startswith('features.') for k in sd) else ''
        with torch.no_grad():
            for name, block in zip(self.names, self.blocks):
                if name == 'pool':
                    continue
                conv = block.layers['conv']
                idx = self._TV_CONV_INDEX[name]
                w, b = sd['%s%d.weight' % (prefix, idx)], sd['%s%d.bias' % (prefix, idx)]
                if tuple(w.shape) != tuple(conv.weight.shape):
                    raise ValueError('VGG-19 checkpoint: %s%d.weight has shape %s, expected %s'
                                     % (prefix, idx, tuple(w.shape), tuple(conv.weight.shape)))
                conv.weight.copy_(w)     # in place: advances the version counter the packed-filter cache keys on
                conv.bias.copy_(b)
        self.pretrained = True

    def forward(self, x):
        out = {}
        for name, block in zip(self.names, self.blocks):
            if name == 'pool':
                x = NF.max_pool2x2(x)
            else:
                x = block.forward_native(x)
                if name in self.wanted:
                    out[name] = x
        return out


_IMAGENET_STATS = {}


def apply_imagenet_normalization(x):
    """[-1, 1] images -> ImageNet-normalised.  ref: imaginaire/utils/misc.py:221-240"""
    key = (x.device, x.dtype)
    if key not in _IMAGENET_STATS:
        # built once per device: creating them from Python lists every call is a pageable
        # host-to-device copy that stalls the launch queue
        _IMAGENET_STATS[key] = (torch.tensor([0.485, 0.456, 0.406], dtype=x.dtype, device=x.device).view(1, 3, 1, 1),
                                torch.tensor([0.229, 0.224, 0.225], dtype=x.dtype, device=x.device).view(1, 3, 1, 1))
    mean, std = _IMAGENET_STATS[key]
    x = (x[:, :3] + 1) / 2
    return (x - mean) / std


class PerceptualLoss(nn.Module):
    def __init__(self, cfg=None, network='vgg19', layers='relu_4_1', weights=None, criterion='l1',
                 resize=False, resize_mode='bilinear', instance_normalized=False, num_scales=1):
        super().__init__()
        if isinstance(layers, str):
            layers = [layers]
        if weights is None:
            weights = [1.0] * len(layers)
        elif isinstance(weights, (float, int)):
            weights = [weights]
        assert len(layers) == len(weights)
        if network != 'vgg19':
            raise NotImplementedError('only the vgg19 perceptual network is provided')
        self.model = _VGG19Features(layers)
        path = None
        if cfg is not None and hasattr(cfg, 'trainer') and hasattr(cfg.trainer, 'perceptual_loss'):
            path = getattr(cfg.trainer.perceptual_loss, 'weights_path', None)
        path = path or os.environ.get('IMX_VGG19_WEIGHTS')
        if path:
            self.model.load_torchvision_state_dict(torch.load(path, map_location='cpu', weights_only=True))
        self._warned = False
        self.layers, self.weights = list(layers), list(weights)
        self.criterion = nn.L1Loss() if criterion == 'l1' else nn.MSELoss()
        self.resize, self.resize_mode = resize, resize_mode
        self.instance_normalized = instance_normalized
        self.num_scales = num_scales

    def forward(self, inp, target):
        self.model.eval()
        if not self.model.pretrained and not self._warned:
            self._warned = True
            warnings.warn('PerceptualLoss: VGG-19 runs with randomly initialised filters (no weights_path / '
                          'IMX_VGG19_WEIGHTS); this is not the reference\'s ImageNet perceptual loss')
        inp, target = apply_imagenet_normalization(inp), apply_imagenet_normalization(target)
        if self.resize:
            inp = F.interpolate(inp, mode=self.resize_mode, size=(224, 224), align_corners=False)
            target = F.interpolate(target, mode=self.resize_mode, size=(224, 224), align_corners=False)
        loss = 0
        for scale in range(self.num_scales):
            feat_in = self.model(inp)
            with torch.no_grad():
                feat_tg = self.model(target)
            for layer, weight in zip(self.layers, self.weights):
                if isinstance(self.criterion, nn.L1Loss) and not self.instance_normalized:
                    # bf16 feature maps: one fused kernel each way instead of ten ATen passes
                    loss = loss + weight * NF.l1_loss_mean(feat_in[layer], feat_tg[layer].detach())
                    continue
                a, b = feat_in[layer].float(), feat_tg[layer].detach().float()
                if self.instance_normalized:
                    a, b = F.instance_norm(a), F.instance_norm(b)
                loss = loss + weight * self.criterion(a, b)
            if scale != self.num_scales - 1:
                inp = F.interpolate(inp, mode=self.resize_mode, scale_factor=0.5, align_corners=False,
                                    recompute_scale_factor=True)
                target = F.interpolate(target, mode=self.resize_mode, scale_factor=0.5, align_corners=False,
                                       recompute_scale_factor=True)
        return loss.float()
