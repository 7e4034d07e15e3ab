"""Losses of the SPADE training step.  These stay plain PyTorch (they are outside the hot
path); only the VGG feature extractor of the perceptual loss is
built from the native Conv2dBlock so that the whole timed step runs on the sm_100a kernels.

ref: imaginaire/losses/gan.py:12-104 (hinge), feature_matching.py:15-60, kl.py,
     perceptual.py:15-191
"""
from .gan import GANLoss
from .feature_matching import FeatureMatchingLoss
from .kl import GaussianKLLoss
from .perceptual import PerceptualLoss

__all__ = ['GANLoss', 'FeatureMatchingLoss', 'GaussianKLLoss', 'PerceptualLoss']
