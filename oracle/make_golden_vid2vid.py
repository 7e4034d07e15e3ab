"""Generates tests/golden/flow_generator_a.npz from the REFERENCE's FlowGenerator + resample() on CPU.

TEST INFRASTRUCTURE ONLY (build container only, like oracle/make_golden.py).

    python -m oracle.make_golden_vid2vid

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


if __name__ == '__main__':
    main()
