"""KL divergence of the style code to N(0, I).  ref: imaginaire/losses/kl.py"""
import torch
from torch import nn


class GaussianKLLoss(nn.Module):
    def forward(self, mu, logvar=None):
        if logvar is None:
            logvar = torch.zeros_like(mu)
        return -0.5 * torch.sum(1 + logvar - mu.pow(2) - logvar.exp())
