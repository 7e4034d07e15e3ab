"""SPADE trainer (generator + discriminator update).
ref: imaginaire/trainers/spade.py:23-187 (constructor, losses, _start_of_iteration, gen_forward, dis_forward),
     imaginaire/trainers/base.py:594-666 (gen_update / dis_update, in trainers/base.py here)

Same constructor as the reference (``Trainer(cfg, net_G, net_D, opt_G, opt_D, sch_G, sch_D, train_data_loader,
val_data_loader)``), same loss names, weights and composition.  The generator runs three times per iteration
exactly as in the reference: once under no_grad for the discriminator update (the fused no-HBM gamma/beta path),
once with autograd for its own update, and once backward.  A network that is frozen in a phase is called through
``.module`` instead of its data-parallel wrapper (same computation, no gradient exchange to set up).
"""
import math

import torch
import torch.nn.functional as F

from ..losses import FeatureMatchingLoss, GANLoss, GaussianKLLoss, PerceptualLoss
from .base import BaseTrainer


class Trainer(BaseTrainer):
    def __init__(self, cfg, net_G, net_D, opt_G, opt_D, sch_G=None, sch_D=None, train_data_loader=(),
                 val_data_loader=None):
        super().__init__(cfg, net_G, net_D, opt_G, opt_D, sch_G, sch_D, train_data_loader, val_data_loader)
        self.video_mode = getattr(cfg.data, 'type', '') == 'imaginaire.datasets.paired_videos'

    def _init_loss(self, cfg):
        t = cfg.trainer
        self.criteria['GAN'] = GANLoss(t.gan_mode)
        self.weights['GAN'] = t.loss_weight.gan
        if hasattr(t, 'perceptual_loss'):
            self.criteria['Perceptual'] = PerceptualLoss(cfg=cfg, network=t.perceptual_loss.mode,
                                                         layers=t.perceptual_loss.layers,
                                                         weights=t.perceptual_loss.weights)
            self.weights['Perceptual'] = t.loss_weight.perceptual
        self.criteria['FeatureMatching'] = FeatureMatchingLoss()
        self.weights['FeatureMatching'] = t.loss_weight.feature_matching
        self.criteria['GaussianKL'] = GaussianKLLoss()
        self.weights['GaussianKL'] = t.loss_weight.kl

    def _start_of_iteration(self, data, current_iteration):
        """Video batches are folded into the channel axis, then labels / images are resized to a multiple of the
        generator's base (spade.py:99-130, 297-313)."""
        if data['label'].dim() == 5:
            label_image_raw = data['images'][:, 0:-1]
            label_image = label_image_raw.reshape(label_image_raw.size(0), -1, label_image_raw.size(3),
                                                  label_image_raw.size(4))
            images = data['images'][:, -1]
            label_label = data['label'].reshape(data['label'].size(0), -1, data['label'].size(3),
                                                data['label'].size(4))
            data['label'] = torch.cat([label_label, label_image], 1)
            data['images'] = images
        return self._resize_data(data)

    def _resize_data(self, data):
        base = getattr(self.net_G, 'base', 32)
        sy = math.floor(data['label'].size(2) * 1.0 // base) * base
        sx = math.floor(data['label'].size(3) * 1.0 // base) * base
        if (sy, sx) != tuple(data['label'].shape[2:]):
            data['label'] = F.interpolate(data['label'], size=[sy, sx], mode='nearest')
            if 'images' in data:
                data['images'] = F.interpolate(data['images'], size=[sy, sx], mode='bicubic')
        return data

    # ----------------------------------------------------------- generator
    def gen_forward(self, data):
        net_G_output = self.net_G(data)
        net_D_output = self.net_D.module(data, net_G_output)      # frozen here: the wrapper has nothing to do
        losses = self.gen_losses
        losses['GAN'] = self.criteria['GAN'](self._get_outputs(net_D_output, real=False), True, dis_update=False)
        losses['FeatureMatching'] = self.criteria['FeatureMatching'](net_D_output['fake_features'],
                                                                     net_D_output['real_features'])
        if self.net_G_module.use_style_encoder:
            losses['GaussianKL'] = self.criteria['GaussianKL'](net_G_output['mu'], net_G_output['logvar'])
        else:
            losses['GaussianKL'] = losses['GAN'].new_zeros(())
        if 'Perceptual' in self.criteria:
            losses['Perceptual'] = self.criteria['Perceptual'](net_G_output['fake_images'], data['images'])
        total = sum(losses[k] * self.weights[k] for k in self.criteria)
        losses['total'] = total
        return total

    # -------------------------------------------------------- discriminator
    def dis_forward(self, data):
        with torch.no_grad():
            net_G_output = self.net_G.module(data)      # no_grad: nothing for the data-parallel wrapper to do
            net_G_output['fake_images'] = net_G_output['fake_images'].detach()
        net_D_output = self.net_D(data, net_G_output)
        fake = self.criteria['GAN'](self._get_outputs(net_D_output, real=False), False, dis_update=True)
        true = self.criteria['GAN'](self._get_outputs(net_D_output, real=True), True, dis_update=True)
        self.dis_losses.update({'GAN/fake': fake, 'GAN/true': true, 'GAN': fake + true})
        total = self.dis_losses['GAN'] * self.weights['GAN']
        self.dis_losses['total'] = total
        return total

    def step(self, data):
        """One training iteration in the reference's order (train.py:76-79): dis_update, then gen_update.
        Returns the two total losses (detached)."""
        self.dis_update(data)
        self.gen_update(data)
        return self.gen_losses['total'], self.dis_losses['total']
