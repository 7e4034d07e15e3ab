"""Activation factory.  ref: imaginaire/layers/nonlinearity.py:8-37

The returned modules are ordinary torch modules (so blocks keep the reference's
``layers.nonlinearity`` entry); inside a Conv2dBlock the hot-path activations
(LeakyReLU(0.2), ReLU) are not run as separate kernels but folded into the
neighbouring native kernel (see conv.py).
"""
from torch import nn


def get_nonlinearity_layer(nonlinearity_type, inplace):
    kind = nonlinearity_type or ''
    if kind in ('', 'none'):
        return None
    if kind == 'relu':
        return nn.ReLU(inplace=inplace)
    if kind == 'leakyrelu':
        return nn.LeakyReLU(0.2, inplace=inplace)   # slope 0.2, nonlinearity.py:22
    if kind == 'prelu':
        return nn.PReLU()
    if kind == 'tanh':
        return nn.Tanh()
    if kind == 'sigmoid':
        return nn.Sigmoid()
    if kind.startswith('softmax'):
        dim = int(kind.split(',')[1]) if ',' in kind else 1
        return nn.Softmax(dim=dim)
    raise ValueError('Nonlinearity %s is not recognized' % nonlinearity_type)
