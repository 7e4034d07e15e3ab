"""Channel-count helpers the generator constructor needs.

ref: imaginaire/utils/data.py:436-483 (channel numbers), :498-520 (get_crop_h_w)
"""


def get_paired_input_image_channel_number(data_cfg):
    total = 0
    for data_type in data_cfg.input_types:
        for name in data_type:
            if name in data_cfg.input_image:
                total += data_type[name].num_channels
    return total


def get_paired_input_label_channel_number(data_cfg, video=False):
    if not hasattr(data_cfg, 'input_labels'):
        return 0
    total = 0
    for data_type in data_cfg.input_types:
        for name in data_type:
            if name in data_cfg.input_labels:
                total += data_type[name].num_channels
                if getattr(data_type[name], 'use_dont_care', False):
                    total += 1      # the extra "don't care" class channel
    if video:
        steps = getattr(data_cfg.train, 'initial_sequence_length', None)
        total = total * steps + get_paired_input_image_channel_number(data_cfg) * (steps - 1)
    return total


def get_crop_h_w(augmentation):
    for key in augmentation.keys():
        if 'crop_h_w' in key:
            crop_h, crop_w = (int(v) for v in str(augmentation[key]).split(','))
            return crop_h, crop_w
    raise AttributeError('no *crop_h_w entry in the augmentation config')
