"""YAML configuration loader compatible with the reference's config files.

ref: imaginaire/config.py:16-213 (AttrDict / Config / recursive_update)

Only the behaviour the hot path relies on is reproduced: nested dicts become
attribute-accessible dicts (``cfg.gen.num_filters``, ``vars(cfg.gen.x)``,
``cfg.data.train.augmentations['random_crop_h_w']``), lists of dicts become lists
of such objects, and a handful of top-level defaults exist before the file is
overlaid, so the reference's ``configs/**.yaml`` load unchanged.
"""
import collections.abc
import re

import yaml


class AttrDict(dict):
    """dict whose keys are also attributes (same object: ``__dict__`` is the dict)."""

    def __init__(self, *args, **kwargs):
        super().__init__(*args, **kwargs)
        self.__dict__ = self
        for key, value in list(self.items()):
            self[key] = _wrap(value)

    def yaml(self):
        out = {}
        for key, value in self.items():
            if isinstance(value, AttrDict):
                out[key] = value.yaml()
            elif isinstance(value, list) and value and isinstance(value[0], AttrDict):
                out[key] = [v.yaml() for v in value]
            else:
                out[key] = value
        return out

    def __repr__(self):
        lines = []
        for key, value in self.items():
            if isinstance(value, AttrDict):
                lines.append('%s:' % key)
                lines += ['    ' + ln for ln in repr(value).split('\n')]
            elif isinstance(value, list) and value and isinstance(value[0], AttrDict):
                lines.append('%s:' % key)
                for item in value:
                    lines += ['    ' + ln for ln in repr(item).split('\n')]
            else:
                lines.append('%s: %s' % (key, value))
        return '\n'.join(lines)


def _wrap(value):
    if isinstance(value, AttrDict):
        return value
    if isinstance(value, dict):
        return AttrDict(value)
    if isinstance(value, (list, tuple)) and len(value) > 0 and isinstance(value[0], dict):
        return [AttrDict(v) for v in value]
    return value


def recursive_update(dst, src):
    for key, value in src.items():
        if isinstance(value, collections.abc.Mapping):
            cur = dst.get(key)
            dst[key] = recursive_update(cur if isinstance(cur, AttrDict) else AttrDict(), value)
        else:
            dst[key] = _wrap(value)
    return dst


def _yaml_loader():
    # YAML 1.1 does not read "1e-4" as a float; the reference patches the resolver the same way.
    loader = yaml.SafeLoader
    loader.add_implicit_resolver(
        'tag:yaml.org,2002:float',
        re.compile(r'''^(?:[-+]?(?:[0-9][0-9_]*)\.[0-9_]*(?:[eE][-+]?[0-9]+)?
                        |[-+]?(?:[0-9][0-9_]*)(?:[eE][-+]?[0-9]+)
                        |\.[0-9_]+(?:[eE][-+][0-9]+)?
                        |[-+]?[0-9][0-9_]*(?::[0-5]?[0-9])+\.[0-9_]*
                        |[-+]?\.(?:inf|Inf|INF)
                        |\.(?:nan|NaN|NAN))$''', re.X),
        list('-+0123456789.'))
    return loader


class Config(AttrDict):
    """``Config(filename)``: defaults overlaid with the YAML file."""

    def __init__(self, filename=None, verbose=False):
        super().__init__()
        # defaults the hot path reads (imaginaire/config.py:77-150)
        self.snapshot_save_iter = 9999999999
        self.snapshot_save_epoch = 9999999999
        self.max_epoch = 9999999999
        self.max_iter = 9999999999
        self.logging_iter = 100
        self.speed_benchmark = False
        self.trainer = AttrDict(
            model_average=False, model_average_beta=0.9999, model_average_start_iteration=1000,
            model_average_batch_norm_estimation_iteration=30, model_average_remove_sn=True,
            image_to_tensorboard=False, hparam_to_tensorboard=False, distributed_data_parallel='pytorch',
            delay_allreduce=True, gan_relativistic=False, gen_step=1, dis_step=1)
        self.gen = AttrDict(type='imaginaire.generators.dummy')
        self.dis = AttrDict(type='imaginaire.discriminators.dummy')
        self.gen_opt = AttrDict(type='adam', fused_opt=False, lr=0.0001, adam_beta1=0.0, adam_beta2=0.999,
                                eps=1e-8, lr_policy=AttrDict(iteration_mode=False, type='step',
                                                             step_size=9999999999, gamma=1))
        self.dis_opt = AttrDict(type='adam', fused_opt=False, lr=0.0001, adam_beta1=0.0, adam_beta2=0.999,
                                eps=1e-8, lr_policy=AttrDict(iteration_mode=False, type='step',
                                                             step_size=9999999999, gamma=1))
        self.data = AttrDict(name='dummy', type='imaginaire.datasets.images', num_workers=0)
        self.test_data = AttrDict(name='dummy', type='imaginaire.datasets.images', num_workers=0,
                                  test=AttrDict(is_lmdb=False, roots='', batch_size=1))
        self.cudnn = AttrDict(deterministic=False, benchmark=True)
        self.pretrained_weight = ''
        self.inference_args = AttrDict()
        if filename is not None:
            with open(filename, 'r') as f:
                cfg_dict = yaml.load(f, Loader=_yaml_loader())
            recursive_update(self, cfg_dict)
            if 'common' in cfg_dict:
                self.common = AttrDict(**cfg_dict['common'])
                self.gen.common = self.common
                self.dis.common = self.common
        if verbose:
            print(' imaginaire config '.center(80, '-'))
            print(repr(self))
            print(''.center(80, '-'))
