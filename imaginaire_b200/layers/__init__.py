"""Layer blocks of the SPADE hot path (same names as imaginaire.layers)."""
from .conv import Conv2d, Conv2dBlock, LinearBlock
from .residual import Res2dBlock
from .activation_norm import AdaptiveNorm, SpatiallyAdaptiveNorm, get_activation_norm_layer
from .nonlinearity import get_nonlinearity_layer
from .weight_norm import get_weight_norm_layer
from .misc import ApplyNoise

__all__ = ['Conv2d', 'Conv2dBlock', 'LinearBlock', 'Res2dBlock', 'AdaptiveNorm', 'SpatiallyAdaptiveNorm',
           'get_activation_norm_layer', 'get_nonlinearity_layer', 'get_weight_norm_layer', 'ApplyNoise']
