"""SPADE training step (generator + discriminator update).
ref: imaginaire/trainers/base.py:594-666 (gen_update / dis_update),
     imaginaire/trainers/spade.py:60-187 (losses, gen_forward, dis_forward)

Only what one optimisation step needs (no logging, checkpointing, FID).  The generator
runs three times per iteration exactly as in the reference: once under no_grad for the
discriminator update (the fused no-HBM gamma/beta path), once with autograd for its own
update, and once backward.
"""
import torch

from .. import grad_sync
from ..losses import FeatureMatchingLoss, GANLoss, GaussianKLLoss, PerceptualLoss
from ..utils.model_average import requires_grad


class Trainer:
    def __init__(self, cfg, net_G, net_D, opt_G, opt_D):
        self.cfg = cfg
        self.net_G, self.net_D, self.opt_G, self.opt_D = net_G, net_D, opt_G, opt_D
        self.net_G_module = net_G.module
        if hasattr(self.net_G_module, 'module') and hasattr(self.net_G_module, 'averaged_model'):
            self.net_G_module = self.net_G_module.module      # inside ModelAverage
        t = cfg.trainer
        self.criteria, self.weights = {}, {}
        self.criteria['GAN'] = GANLoss(t.gan_mode)
        self.weights['GAN'] = t.loss_weight.gan
        if hasattr(t, 'perceptual_loss'):
            dev = next(net_G.parameters()).device
            self.criteria['Perceptual'] = PerceptualLoss(cfg=cfg, network=t.perceptual_loss.mode,
                                                         layers=t.perceptual_loss.layers,
                                                         weights=t.perceptual_loss.weights).to(dev)
            self.weights['Perceptual'] = t.loss_weight.perceptual
        self.criteria['FeatureMatching'] = FeatureMatchingLoss()
        self.weights['FeatureMatching'] = t.loss_weight.feature_matching
        self.criteria['GaussianKL'] = GaussianKLLoss()
        self.weights['GaussianKL'] = t.loss_weight.kl
        self.gen_losses, self.dis_losses = {}, {}

    # ----------------------------------------------------------- generator
    def gen_forward(self, data):
        net_G_output = self.net_G(data)
        # the discriminator is frozen here: no gradient exchange to set up, so the data-parallel
        # wrapper is bypassed (same computation)
        net_D_output = self.net_D.module(data, net_G_output)
        losses = self.gen_losses
        losses['GAN'] = self.criteria['GAN'](net_D_output['fake_outputs'], True, dis_update=False)
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

    def gen_update(self, data):
        self.opt_G.zero_grad(set_to_none=True)
        requires_grad(self.net_G_module, True)
        requires_grad(self.net_D, False)
        total = self.gen_forward(data)
        total.backward()
        grad_sync.finish_all()      # no-op unless the peer-memory gradient path is active
        self.opt_G.step()
        if self.cfg.trainer.model_average:
            self.net_G.module.update_average()
        return total.detach()

    # -------------------------------------------------------- discriminator
    def dis_forward(self, data):
        with torch.no_grad():
            net_G_output = self.net_G.module(data)      # no_grad: nothing for the DDP wrapper to do
            net_G_output['fake_images'] = net_G_output['fake_images'].detach()
        net_D_output = self.net_D(data, net_G_output)
        fake = self.criteria['GAN'](net_D_output['fake_outputs'], False, dis_update=True)
        true = self.criteria['GAN'](net_D_output['real_outputs'], True, dis_update=True)
        self.dis_losses.update({'GAN/fake': fake, 'GAN/true': true, 'GAN': fake + true})
        total = self.dis_losses['GAN'] * self.weights['GAN']
        self.dis_losses['total'] = total
        return total

    def dis_update(self, data):
        self.opt_D.zero_grad(set_to_none=True)
        requires_grad(self.net_G_module, False)
        requires_grad(self.net_D, True)
        total = self.dis_forward(data)
        total.backward()
        grad_sync.finish_all()
        self.opt_D.step()
        return total.detach()

    def step(self, data):
        """One training iteration in the reference's order (train.py:76-79)."""
        d = self.dis_update(data)
        g = self.gen_update(data)
        return g, d
