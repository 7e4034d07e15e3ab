"""Discriminator feature matching.  ref: imaginaire/losses/feature_matching.py:15-60"""
from torch import nn

from .. import functional as NF


class FeatureMatchingLoss(nn.Module):
    def __init__(self, criterion='l1'):
        super().__init__()
        if criterion == 'l1':
            self.criterion = nn.L1Loss()
        elif criterion in ('l2', 'mse'):
            self.criterion = nn.MSELoss()
        else:
            raise ValueError('Criterion %s is not recognized' % criterion)

    def forward(self, fake_features, real_features):
        num_d = len(fake_features)
        loss = 0.0
        for fake_d, real_d in zip(fake_features, real_features):
            for fake, real in zip(fake_d, real_d):
                if isinstance(self.criterion, nn.L1Loss):
                    loss = loss + NF.l1_loss_mean(fake, real.detach()) / num_d
                else:
                    loss = loss + self.criterion(fake.float(), real.detach().float()) / num_d
        return loss
