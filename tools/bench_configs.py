"""Measurements for the other BASELINE.json configurations; developer tool.

    python tools/bench_configs.py config1   # unit-test SPADE generator forward, batch 1 (GPU + CPU oracle)
    python tools/bench_configs.py config4   # flow-warp microbenchmark, 512x1024, B in {1,8,32} x C in {3,4,64}
    python tools/bench_configs.py config5   # C3 generator at 512x1024, batch 4, forward only
    python tools/bench_configs.py all

One JSON line per measurement.  bench.py (config 3, the training step) stays the driver's contract.
GPU times: CUDA events around REPS back-to-back calls after warm-up; inputs larger than L2 or
rotated between calls where it matters.
"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from imaginaire_b200.config import Config

_PEAKS = bench.load_peaks()
HBM_PEAK = _PEAKS['hbm']
DEV = 'cuda:0'


def gpu_time(fn, reps, warmup=2):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(reps):
        fn()
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) / reps * 1e-3


def config1():
    from imaginaire_b200.generators.spade import Generator
    from imaginaire_b200.utils.init_weight import weights_init
    cfg = Config(os.path.join(bench.ROOT, 'configs', 'spade_unit_test.yaml'))
    torch.manual_seed(0)
    net = Generator(cfg.gen, cfg.data)
    net.apply(weights_init('xavier', 0.02))
    net = net.to(DEV).train()
    data = {k: v.to(DEV) for k, v in bench.synthetic_batch(cfg, 1, 0).items()}
    with torch.no_grad():
        t = gpu_time(lambda: net(data), 10, 3)
    torch.set_num_threads(os.cpu_count() or 1)
    cpu = min(bench.cpu_reference_step(cfg, 1, 0)[0] for _ in range(3))
    print(json.dumps({'config': 'config1: configs/unit_test/spade.yaml generator forward, batch 1, 256x256, 185-channel label',
                      'gpu_ms': t * 1e3, 'gpu_img_s': 1.0 / t, 'cpu_oracle_ms': cpu * 1e3, 'cpu_cores': os.cpu_count(),
                      'gflop_per_img': 420.0, 'gpu_tflops': 0.420 / t, 'mode': 'train-mode norms, no_grad'}))


def synthetic_flow(b, h, w, gen, grid=(32, 64)):
    """bilinear upsample of N(0, 8 px) drawn at 32x64 (``grid``: coarser grids
    give smoother fields of the same amplitude)."""
    low = torch.randn(b, 2, grid[0], grid[1], generator=gen) * 8.0
    return torch.nn.functional.interpolate(low, size=(h, w), mode='bilinear', align_corners=False).contiguous()


def config4_smoothness():
    """The same warps as config4 (B=8, C in {3, 64}) under smoother flow fields: N(0, 8 px) drawn on
    coarser grids, and a constant translation.  Shows how much of the distance to the HBM peak is
    the gather pattern of the SURVEY flow (|grad flow| ~0.7 px/px) rather than the kernel."""
    from imaginaire_b200 import ops
    h, w, b = 512, 1024, 8
    g = torch.Generator().manual_seed(4)
    for c in (3, 64):
        img = torch.randn(b, c, h, w, generator=g).to(DEV)
        go = torch.randn(b, c, h, w, generator=g).to(DEV)
        fwd_bytes = (2 * c + 2) * b * h * w * 4
        bwd_bytes = (3 * c + 4) * b * h * w * 4
        for label, grid in (('N(0,8px) at 32x64 (SURVEY)', (32, 64)), ('N(0,8px) at 8x16', (8, 16)), ('N(0,8px) at 2x4', (2, 4)),
                            ('constant (3.3, -2.6) px', None)):
            if grid is None:
                flow = torch.empty(b, 2, h, w)
                flow[:, 0] = 3.3
                flow[:, 1] = -2.6
            else:
                flow = synthetic_flow(b, h, w, g, grid)
            gx = (flow[:, :, :, 1:] - flow[:, :, :, :-1]).abs().mean().item()
            flow = flow.to(DEV)
            reps = 20 if c == 3 else 5
            tf = gpu_time(lambda: ops.resample2d_fwd(img, flow), reps)
            tb = gpu_time(lambda: ops.resample2d_bwd(img, flow, go), reps)
            tg = gpu_time(lambda: ops.gridsample_warp_fwd(img, flow), reps)
            print(json.dumps({'config': 'config4 warp smoothness B=%d C=%d 512x1024, flow %s' % (b, c, label),
                              'mean_abs_dflow_dx': gx,
                              'resample2d_fwd_us': tf * 1e6, 'resample2d_fwd_frac_hbm': fwd_bytes / tf / 1e9 / HBM_PEAK,
                              'resample_fwd_us': tg * 1e6, 'resample_fwd_frac_hbm': fwd_bytes / tg / 1e9 / HBM_PEAK,
                              'resample2d_bwd_us': tb * 1e6, 'resample2d_bwd_frac_hbm': bwd_bytes / tb / 1e9 / HBM_PEAK}))
        del img, go


def config4():
    from imaginaire_b200 import ops
    h, w = 512, 1024
    g = torch.Generator().manual_seed(4)
    for b in (1, 8, 32):
        for c in (3, 4, 64):
            if b * c * h * w * 4 > 6e9:
                continue
            img = torch.randn(b, c, h, w, generator=g).to(DEV)
            flow = synthetic_flow(b, h, w, g).to(DEV)
            go = torch.randn(b, c, h, w, generator=g).to(DEV)
            fwd_bytes = (2 * c + 2) * b * h * w * 4
            bwd_bytes = (3 * c + 4) * b * h * w * 4
            reps = 20 if b * c <= 64 else 5
            for name, f_fwd, f_bwd in (('resample2d', lambda: ops.resample2d_fwd(img, flow),
                                        lambda: ops.resample2d_bwd(img, flow, go)),
                                       ('resample (grid_sample form)', lambda: ops.gridsample_warp_fwd(img, flow),
                                        lambda: ops.gridsample_warp_bwd(img, flow, go))):
                tf = gpu_time(f_fwd, reps)
                tb = gpu_time(f_bwd, reps)
                print(json.dumps({'config': 'config4 warp %s B=%d C=%d 512x1024' % (name, b, c),
                                  'fwd_us': tf * 1e6, 'fwd_GBps': fwd_bytes / tf / 1e9, 'fwd_frac_hbm': fwd_bytes / tf / 1e9 / HBM_PEAK,
                                  'bwd_us': tb * 1e6, 'bwd_GBps': bwd_bytes / tb / 1e9, 'bwd_frac_hbm': bwd_bytes / tb / 1e9 / HBM_PEAK,
                                  'note': 'B=1: latency, not bandwidth' if b == 1 else ''}))
            del img, flow, go
    # CPU restatement of resample2d beside it (C oracle, single thread), B=1 C=3
    from oracle import warp_oracle as wo
    import numpy as np
    img = np.random.default_rng(0).standard_normal((1, 3, h, w), dtype=np.float32)
    flow = synthetic_flow(1, h, w, g).numpy()
    t0 = time.perf_counter()
    wo.resample2d_forward(img, flow)
    t = time.perf_counter() - t0
    print(json.dumps({'config': 'config4 cpu oracle resample2d fwd B=1 C=3 512x1024', 'cpu_ms': t * 1e3, 'cores': 1,
                      'GBps': (2 * 3 + 2) * h * w * 4 / t / 1e9}))


def config4_flow_branch():
    """vid2vid flow branch at 512x1024 (Cityscapes-shaped: 35 label channels, 3 frames): FlowGenerator
    (defaults of configs/projects/vid2vid/cityscapes/ampO1.yaml: 32 filters, 5 downsamples, 6 res blocks) ->
    flow, mask -> warp of the previous frame -> composite; forward + backward, batch 1 and 4."""
    import types
    from imaginaire_b200.generators.vid2vid import FlowGenerator, warp_prev_frame
    from imaginaire_b200.utils.init_weight import weights_init
    ns = types.SimpleNamespace
    data_cfg = ns(input_types=[{'images': ns(num_channels=3)}, {'seg_maps': ns(num_channels=35)}],
                  input_image=['images'], input_labels=['seg_maps'], num_frames_G=3)
    torch.manual_seed(0)
    net = FlowGenerator(ns(), data_cfg)
    net.apply(weights_init('xavier', 0.02))
    net = net.to(DEV).train()
    h, w = 512, 1024
    for b in (1, 4):
        g = torch.Generator().manual_seed(5)
        idx = torch.randint(0, 35, (b, 3, h // 8, w // 8), generator=g).repeat_interleave(8, 2).repeat_interleave(8, 3)
        label = torch.zeros(b, 3, 35, h, w).scatter_(2, idx.unsqueeze(2), 1.0).view(b, 105, h, w).to(DEV)
        img_prev = (torch.rand(b, 6, h, w, generator=g) * 2 - 1).to(DEV)
        raw = (torch.rand(b, 3, h, w, generator=g) * 2 - 1).to(DEV)

        def step():
            flow, mask = net(label, img_prev)
            warp = warp_prev_frame(img_prev[:, -3:], flow)
            final = raw * mask + warp * (1 - mask)
            final.square().mean().backward()
        with torch.no_grad():
            tf = gpu_time(lambda: net(label, img_prev), 5, 2)
        t = gpu_time(step, 5, 2)
        print(json.dumps({'config': 'config4 vid2vid flow branch (FlowGenerator + warp + composite) 512x1024 B=%d' % b,
                          'fwd_ms': tf * 1e3, 'fwd_bwd_ms': t * 1e3, 'frames_per_s_train': b / t}))


def config4_generator():
    """vid2vid generator of configs/projects/vid2vid/cityscapes/ampO1.yaml (num_layers 7, 32 filters, instance-norm
    SPADE with 1x1 projections on embedded labels, unet image embedder, multi-SPADE combine), 512x1024, 35 labels,
    temporal frame (previous-frame encoder + flow network + warp + generator), forward and forward+backward."""
    import types
    from imaginaire_b200.generators.vid2vid import Generator
    from imaginaire_b200.utils.init_weight import weights_init
    ns = types.SimpleNamespace
    data_cfg = ns(input_types=[{'images': ns(num_channels=3)}, {'seg_maps': ns(num_channels=35)}],
                  input_image=['images'], input_labels=['seg_maps'], num_frames_G=3,
                  val=ns(augmentations=ns(center_crop_h_w='512, 1024')))
    emb = lambda arch: ns(use_embed=True, arch=arch, num_filters=32, num_downsamples=5, kernel_size=3,
                          weight_norm_type='spectral')
    flow = ns(num_filters=32, max_num_filters=1024, num_downsamples=5, num_res_blocks=6, activation_norm_type='instance',
              weight_norm_type='spectral', flow_output_multiplier=40, generate_raw_output=False,
              multi_spade_combine=ns(num_layers=3, embed=emb('unet')))
    gen_cfg = ns(num_layers=7, num_downsamples_img=4, num_filters=32, max_num_filters=1024, kernel_size=3,
                 activation_norm_type='spatially_adaptive',
                 activation_norm_params=ns(activation_norm_type='instance', num_filters=0, kernel_size=1),
                 weight_norm_type='spectral', style_dims=256, use_segmap_as_input=True, flow=flow, embed=emb('encoderdecoder'))
    torch.manual_seed(0)
    net = Generator(gen_cfg, data_cfg)
    net.apply(weights_init('xavier', 0.02))
    net = net.to(DEV).train()
    h, w, b = 512, 1024, 1
    g = torch.Generator().manual_seed(6)
    idx = torch.randint(0, 35, (b, 3, h // 8, w // 8), generator=g).repeat_interleave(8, 2).repeat_interleave(8, 3)
    labels = torch.zeros(b, 3, 35, h, w).scatter_(2, idx.unsqueeze(2), 1.0).to(DEV)
    imgs = (torch.rand(b, 2, 3, h, w, generator=g) * 2 - 1).to(DEV)
    first = {'label': labels[:, 0], 'prev_labels': None, 'prev_images': None}
    temporal = {'label': labels[:, 2], 'prev_labels': labels[:, :2], 'prev_images': imgs}
    with torch.no_grad():
        t_first = gpu_time(lambda: net(first), 5, 2)
        t_temp = gpu_time(lambda: net(temporal), 5, 2)
    t_train = gpu_time(lambda: net(temporal)['fake_images'].square().mean().backward(), 5, 2)
    print(json.dumps({'config': 'config4 vid2vid generator (cityscapes ampO1 architecture) 512x1024 B=1',
                      'params_M': sum(p.numel() for p in net.parameters()) / 1e6,
                      'first_frame_fwd_ms': t_first * 1e3, 'temporal_frame_fwd_ms': t_temp * 1e3,
                      'temporal_frame_fwd_bwd_ms': t_train * 1e3, 'frames_per_s_inference': 1.0 / t_temp}))


def config5():
    from imaginaire_b200.generators.spade import Generator
    from imaginaire_b200.utils.init_weight import weights_init
    cfg = Config(os.path.join(bench.ROOT, 'configs', 'spade_c3_256x512.yaml'))
    cfg.data.train.augmentations.random_crop_h_w = '512, 1024'
    torch.manual_seed(0)
    net = Generator(cfg.gen, cfg.data)
    net.apply(weights_init('xavier', 0.02))
    net = net.to(DEV)
    batch = 4
    data = {k: v.to(DEV) for k, v in bench.synthetic_batch(cfg, batch, 0).items()}
    for mode in ('train', 'eval'):
        net.train(mode == 'train')
        with torch.no_grad():
            for _ in range(2):
                out = net(data)        # train mode: also warms the running statistics used by eval
            assert tuple(out['fake_images'].shape) == (batch, 3, 512, 1024)
            assert bool(torch.isfinite(out['fake_images']).all())
            t = gpu_time(lambda: net(data), 5, 1)
        print(json.dumps({'config': 'config5: C3 generator, 512x1024, batch %d, forward only, %s-mode norms' % (batch, mode),
                          'ms': t * 1e3, 'img_s': batch / t, 'gflop_per_img': 4745.0, 'tflops': 4.745 * batch / t,
                          'frac_of_sustained_bf16': 4.745 * batch / t / _PEAKS['tf_sustained']}))


if __name__ == '__main__':
    which = sys.argv[1] if len(sys.argv) > 1 else 'all'
    for name, fn in (('config1', config1), ('config4', config4), ('config4', config4_flow_branch), ('config4', config4_generator), ('config5', config5),
                     ('config4_smoothness', config4_smoothness)):
        if which == name or (which == 'all' and name != 'config4_smoothness'):
            fn()
