"""Weight-normalisation factory.  ref: imaginaire/layers/weight_norm.py:71-92

'spectral' uses torch.nn.utils.spectral_norm exactly as the reference does, so the
parameter names (weight_orig / weight_u / weight_v), the one-step power iteration in
training mode and the interaction with ModelAverage are the reference's.  The hook
leaves ``module.weight = weight_orig / sigma``; the native conv then casts and repacks
that tensor into the tcgen05 operand layout in one kernel (ops.pack_conv_weight).
"""
import functools

from torch.nn.utils import spectral_norm, weight_norm


def get_weight_norm_layer(norm_type, **norm_params):
    if norm_type in ('none', ''):
        return lambda x: x
    if norm_type == 'spectral':
        return functools.partial(spectral_norm, **norm_params)
    if norm_type == 'weight':
        return functools.partial(weight_norm, **norm_params)
    if norm_type == 'weight_demod':
        raise NotImplementedError("weight_demod is outside the SPADE hot path (SURVEY.md section 2.1 #4)")
    raise ValueError('Weight norm layer %s is not recognized' % norm_type)
