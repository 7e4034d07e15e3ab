"""Generates tests/golden/flow_generator_a.npz from the REFERENCE's FlowGenerator + resample() on CPU.

TEST INFRASTRUCTURE ONLY (build container only, like oracle/make_golden.py).

    python -m oracle.make_golden_vid2vid              # flow_generator_a.npz
    python -m oracle.make_golden_vid2vid generator    # vid2vid_generator_a.npz

ref: imaginaire/generators/vid2vid.py:390-481 (FlowGenerator), :230 (warp of the previous frame),
     imaginaire/utils/trainer.py:103-106 (weights_init as the trainer applies it)
"""
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)

from oracle import ref_shims  # noqa: E402
from oracle.make_golden import onehot_label, save, sd_arrays  # noqa: E402


def configs():
    ns = types.SimpleNamespace
    data_cfg = ns(input_types=[{'images': ns(num_channels=3)}, {'seg_maps': ns(num_channels=12)}],
                  input_image=['images'], input_labels=['seg_maps'], num_frames_G=3)
    flow_cfg = ns(num_filters=16, max_num_filters=64, num_downsamples=2, kernel_size=3, num_res_blocks=2,
                  flow_output_multiplier=20, activation_norm_type='sync_batch', weight_norm_type='spectral')
    return flow_cfg, data_cfg


def main():
    ref_shims.install()
    from imaginaire.generators.vid2vid import FlowGenerator
    from imaginaire.model_utils.fs_vid2vid import resample
    from imaginaire.utils.init_weight import weights_init
    flow_cfg, data_cfg = configs()
    torch.manual_seed(11)
    net = FlowGenerator(flow_cfg, data_cfg)
    net.apply(weights_init('xavier', 0.02))
    net.train()
    sd_before = {k: v.clone() for k, v in net.state_dict().items()}
    b, h, w = 2, 32, 48
    label = torch.cat([onehot_label(20 + t, b, 12, h, w, blocky=4) for t in range(3)], 1)    # 3 frames of labels
    g = torch.Generator().manual_seed(21)
    img_prev = torch.rand(b, 6, h, w, generator=g) * 2 - 1                                    # 2 previous frames
    with torch.no_grad():
        flow, mask = net(label, img_prev)
        warp = resample(img_prev[:, -3:], flow)
    save('flow_generator_a.npz', label=label, img_prev=img_prev, flow=flow, mask=mask, warp=warp,
         keys=np.array(list(sd_before.keys())), **sd_arrays(sd_before, 'sd.'), **sd_arrays(net.state_dict(), 'sd_after.'))
    print('flow', float(flow.abs().mean()), 'mask', float(mask.mean()))


if __name__ == '__main__' and len(sys.argv) == 1:
    main()


def generator_configs():
    """Small vid2vid generator: 4 up layers at 64x96, instance-norm SPADE with 1x1 projections on the
    embedded labels (configs/projects/vid2vid/cityscapes/ampO1.yaml in miniature), multi-SPADE combine."""
    ns = types.SimpleNamespace
    data_cfg = ns(input_types=[{'images': ns(num_channels=3)}, {'seg_maps': ns(num_channels=12)}],
                  input_image=['images'], input_labels=['seg_maps'], num_frames_G=3,
                  val=ns(augmentations=ns(center_crop_h_w='64, 96')))
    flow_cfg = ns(num_filters=8, max_num_filters=32, num_downsamples=2, kernel_size=3, num_res_blocks=2,
                  flow_output_multiplier=20, activation_norm_type='sync_batch', weight_norm_type='spectral',
                  generate_raw_output=False,
                  multi_spade_combine=ns(num_layers=2, embed=ns(use_embed=True, arch='unet', num_filters=8,
                                                                num_downsamples=3, kernel_size=3,
                                                                weight_norm_type='spectral')))
    gen_cfg = ns(num_layers=4, num_downsamples_img=2, num_filters=8, max_num_filters=64, kernel_size=3,
                 style_dims=32, weight_norm_type='spectral', activation_norm_type='spatially_adaptive',
                 activation_norm_params=ns(activation_norm_type='instance', kernel_size=1,
                                           separate_projection=False, weight_norm_type='spectral',
                                           activation_norm_params=ns(affine=False)),
                 embed=ns(use_embed=True, arch='encoderdecoder', num_filters=8, num_downsamples=3, kernel_size=3,
                          weight_norm_type='spectral'),
                 flow=flow_cfg)
    return gen_cfg, data_cfg


def main_generator():
    ref_shims.install()
    from imaginaire.generators.vid2vid import FlowGenerator, Generator
    from imaginaire.utils.init_weight import weights_init
    # vid2vid.py:185 asks for label.get_device() (-1 on the CPU): hand back the device object instead
    torch.Tensor.get_device = lambda self: self.device
    gen_cfg, data_cfg = generator_configs()
    torch.manual_seed(12)
    net = Generator(gen_cfg, data_cfg)
    # the fork comments out this upstream line of init_temporal_network
    net.flow_network_temp = FlowGenerator(gen_cfg.flow, data_cfg)
    net.apply(weights_init('xavier', 0.02))
    net.train()
    sd_before = {k: v.clone() for k, v in net.state_dict().items()}
    b, h, w = 2, 64, 96
    labels = [onehot_label(40 + t, b, 12, h, w, blocky=8) for t in range(3)]
    g = torch.Generator().manual_seed(41)
    imgs = [torch.rand(b, 3, h, w, generator=g) * 2 - 1 for _ in range(2)]
    with torch.no_grad():
        first = net({'label': labels[0], 'prev_labels': None, 'prev_images': None})
        temporal = net({'label': labels[2], 'prev_labels': torch.stack(labels[:2], 1), 'prev_images': torch.stack(imgs, 1)})
    save('vid2vid_generator_a.npz', labels=torch.stack(labels, 1), prev_images=torch.stack(imgs, 1),
         first_fake=first['fake_images'], temporal_fake=temporal['fake_images'], temporal_flow=temporal['fake_flow_maps'],
         temporal_mask=temporal['fake_occlusion_masks'], temporal_warp=temporal['warped_images'],
         keys=np.array(sorted(sd_before.keys())), **sd_arrays(sd_before, 'sd.'))
    print('first', float(first['fake_images'].abs().mean()), 'temporal', float(temporal['fake_images'].abs().mean()),
          'flow', float(temporal['fake_flow_maps'].abs().mean()))


if __name__ == '__main__' and len(sys.argv) > 1 and sys.argv[1] == 'generator':
    main_generator()
