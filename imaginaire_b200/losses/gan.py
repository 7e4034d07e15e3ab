"""GAN objectives.  ref: imaginaire/losses/gan.py:12-104"""
import torch
import torch.nn.functional as F
from torch import nn


class GANLoss(nn.Module):
    def __init__(self, gan_mode, target_real_label=1.0, target_fake_label=0.0):
        super().__init__()
        self.gan_mode = gan_mode
        self.real_label = target_real_label
        self.fake_label = target_fake_label

    def forward(self, dis_output, t_real, dis_update=True):
        if isinstance(dis_output, (list, tuple)):
            return sum(self.loss(o, t_real, dis_update) for o in dis_output) / len(dis_output)
        return self.loss(dis_output, t_real, dis_update)

    def loss(self, dis_output, t_real, dis_update=True):
        if not dis_update:
            assert t_real, 'The target should be real when updating the generator.'
        x = dis_output.float()
        if self.gan_mode == 'hinge':
            if not dis_update:
                return -x.mean()
            # -mean(min(+-x - 1, 0))
            return F.relu(1.0 - x).mean() if t_real else F.relu(1.0 + x).mean()
        target = torch.full_like(x, self.real_label if t_real else self.fake_label)
        if self.gan_mode == 'non_saturated':
            return F.binary_cross_entropy_with_logits(x, target)
        if self.gan_mode == 'least_square':
            return 0.5 * F.mse_loss(x, target)
        if self.gan_mode == 'wasserstein':
            return -x.mean() if t_real else x.mean()
        raise ValueError('Unexpected gan_mode {}'.format(self.gan_mode))
