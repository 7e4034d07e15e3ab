"""Small helper layers.  ref: imaginaire/layers/misc.py (ApplyNoise only)."""
import torch
from torch import nn


class ApplyNoise(nn.Module):
    """Adds Gaussian noise with a learnable magnitude (not used by the SPADE generator;
    kept so that ``apply_noise=True`` blocks construct with the reference's keys)."""

    def __init__(self):
        super().__init__()
        self.weight = nn.Parameter(torch.zeros(1))

    def forward(self, x, noise=None):
        if noise is None:
            sz = x.size()
            noise = x.new_empty(sz[0], 1, *sz[2:]).normal_()
        return x + self.weight * noise
