"""Weight initialisation applied by the trainer.  ref: imaginaire/utils/init_weight.py:8-61

Note a property of the reference that this keeps: for spectrally-normalised layers
``m.weight`` is the derived tensor W/sigma, so initialising ``m.weight.data`` only
consumes random numbers; ``weight_orig`` keeps torch's default initialisation, and the
bias is still zeroed.
"""
from torch.nn import init


def weights_init(init_type='normal', gain=0.02, bias=None):
    def init_func(m):
        name = m.__class__.__name__
        if not (hasattr(m, 'weight') and any(s in name for s in ('Conv', 'Linear', 'Embedding'))):
            return
        w = m.weight.data
        if init_type == 'normal':
            init.normal_(w, 0.0, gain)
        elif init_type == 'xavier':
            init.xavier_normal_(w, gain=gain)
        elif init_type == 'xavier_uniform':
            init.xavier_uniform_(w, gain=1.0)
        elif init_type == 'kaiming':
            init.kaiming_normal_(w, a=0, mode='fan_in')
        elif init_type == 'orthogonal':
            init.orthogonal_(w, gain=gain)
        elif init_type == 'none':
            m.reset_parameters()
        else:
            raise NotImplementedError('initialization method [%s] is not implemented' % init_type)
        if getattr(m, 'bias', None) is not None:
            if bias is None:
                init.constant_(m.bias.data, 0.0)
            elif getattr(bias, 'type', 'normal') == 'normal':
                init.normal_(m.bias.data, 0.0, getattr(bias, 'gain', 0.5))
            else:
                raise NotImplementedError('bias initialization [%s] is not implemented' % bias.type)
    return init_func
