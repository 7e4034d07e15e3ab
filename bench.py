"""Benchmark of the SPADE hot path on B200 (driver contract: see the task description).

    python bench.py --gpus 1 --steps K --warmup W            # our arm
    python bench.py --impl reference --steps K --warmup W    # reference arm: CPU path on host cores
    torchrun ... bench.py --gpus N ...                       # one rank per GPU, NCCL

Workload (BASELINE.json configs[2], SURVEY.md section 8(d) "config 3"): one SPADE G+D
training iteration of the base128 generator (configs/spade_c3_256x512.yaml) at 256x512,
35-class synthetic one-hot labels, batch 8 per GPU, bf16 storage / fp32 accumulate:
discriminator update (no-grad generator pass + D forward/backward + Adam) then generator
update (G forward, D forward, GAN + feature-matching + KL + VGG perceptual losses, backward,
Adam, EMA).  Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = 'spade_gd_train_images_per_sec_256x512'
UNIT = 'img/s'
CONFIG_NAME = 'spade_c3_256x512.yaml'


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=5)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--config', default=CONFIG_NAME)
    ap.add_argument('--batch', type=int, default=None, help='per-GPU batch override (debugging only)')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--no-hbm-probe', action='store_true', help='skip the norm / warp HBM microbenchmarks')
    ap.add_argument('--no-graph', action='store_true', help='single-GPU run: time the eager loop instead of the CUDA-graph replay')
    ap.add_argument('--no-clocks', action='store_true', help='do not spawn the nvidia-smi clock sampler (ncu runs)')
    ap.add_argument('--one-step', action='store_true',
                    help='profiling aid: run exactly one training iteration (after --warmup) and exit, '
                         'so that an ncu launch list covers one step and nothing else')
    return ap.parse_args()


def load_peaks():
    path = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(path):
        with open(path) as f:
            p = json.load(f)
        return dict(hbm=p['hbm_gbs'], tf_burst=p['bf16_tflops'], tf_sustained=p['bf16_tflops_sustained'],
                    source='MEASURED_PEAKS.json')
    return dict(hbm=6650.0, tf_burst=1590.0, tf_sustained=1400.0, source='fallback (B200_PROFILING.md)')


# --------------------------------------------------------------------- synthetic data

def synthetic_batch(cfg, batch, seed, pinned=False):
    """Labels: one-hot of a blocky randint class map (what real label maps look like,
    SURVEY 8(d)); images: U(-1, 1).  Host tensors, fp32, like the reference's data loader."""
    import torch
    from imaginaire_b200.utils.data import get_crop_h_w, get_paired_input_label_channel_number
    h, w = get_crop_h_w(cfg.data.train.augmentations)
    classes = get_paired_input_label_channel_number(cfg.data)
    g = torch.Generator().manual_seed(seed)
    idx = torch.randint(0, classes, (batch, 1, h // 8, w // 8), generator=g)
    idx = idx.repeat_interleave(8, 2).repeat_interleave(8, 3)
    label = torch.zeros(batch, classes, h, w).scatter_(1, idx, 1.0)
    images = torch.rand(batch, 3, h, w, generator=g) * 2 - 1
    if pinned:
        label, images = label.pin_memory(), images.pin_memory()
    return {'label': label, 'images': images}


# ------------------------------------------------------------------------- clocks

class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons while the timed region runs."""
    FIELDS = ('clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,'
              'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')

    def __init__(self, index):
        self.index, self.samples, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ['nvidia-smi', '-i', str(self.index), '--query-gpu=' + self.FIELDS, '--format=csv,noheader,nounits',
                 '-lms', '200'], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            parts = [p.strip() for p in line.split(',')]
            if len(parts) >= 6:
                self.samples.append(parts)

    def stop(self):
        if self.proc is not None:
            self.proc.terminate()
        sm = sorted(int(s[0]) for s in self.samples if s[0].isdigit())
        mx = [int(s[1]) for s in self.samples if s[1].isdigit()]
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        reasons = sorted({names[i] for s in self.samples for i in range(4) if s[2 + i].lower().startswith('active')})
        return {'sm_mhz': sm[len(sm) // 2] if sm else None, 'sm_max_mhz': max(mx) if mx else None,
                'reasons': reasons, 'samples': len(self.samples)}


# ------------------------------------------------------------- conv roofline probe

class ConvProbe:
    """Times every implicit-GEMM convolution launch of one training iteration with CUDA events
    on the launching stream and adds up its algorithmic FLOPs (2*N*Ho*Wo*Cout*Cin*kh*kw)."""

    def __init__(self):
        self.records = []

    def install(self):
        import torch
        from imaginaire_b200 import ops
        self._orig = (ops.conv2d_fprop, ops.conv2d_wgrad, ops.spade_gb_conv_modulate, ops.conv2d_fprop_window)
        probe = self

        def timed(fn, flops_fn, tag):
            def wrapper(*a, **k):
                s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                s.record()
                out = fn(*a, **k)
                e.record()
                probe.records.append((tag, flops_fn(out, a, k), s, e, probe.describe(tag, out, a, k)))
                return out
            return wrapper

        # (result, positional arguments, keyword arguments) of the wrapped call
        def fprop_flops(out, a, k):
            x, wpk, cout, kh, kw = a[:5]
            # algorithmic: the logical channel counts (a 35-channel label map is stored with 40 channels, an image
            # with 8; the padding is the kernel's business, not work the reference does)
            n, _, ho, wo = out.shape
            return 2.0 * n * ho * wo * cout * getattr(wpk, '_imx_inner', x.shape[1]) * kh * kw

        def wgrad_flops(out, a, k):
            x, dy, cin, cout, kh, kw = a[:6]
            n, _, ho, wo = dy.shape
            return 2.0 * n * ho * wo * cout * cin * kh * kw

        def gb_flops(out, a, k):
            hidden, c, kh, kw = a[0], a[9], a[10], a[11]
            n, chp, h, wd = hidden.shape
            return 2.0 * n * h * wd * (2 * c) * chp * kh * kw

        def window_flops(out, a, k):      # one parity class of a stride-2 data gradient
            x, wpk, cout, kh, kw, _, _, ho, wo = a[:9]
            return 2.0 * x.shape[0] * ho * wo * cout * getattr(wpk, '_imx_inner', x.shape[1]) * kh * kw

        ops.conv2d_fprop_window = timed(self._orig[3], window_flops, 'fprop')
        ops.conv2d_fprop = timed(self._orig[0], fprop_flops, 'fprop')
        ops.conv2d_wgrad = timed(self._orig[1], wgrad_flops, 'wgrad')
        ops.spade_gb_conv_modulate = timed(self._orig[2], gb_flops, 'spade_gb')

    @staticmethod
    def describe(tag, out, a, k):
        if tag == 'fprop' and len(a) > 12:
            return 'x%s k%dx%d window %dx%d of %s (stride-2 dgrad class)' % (tuple(a[0].shape), a[3], a[4], a[7], a[8], tuple(a[9].shape))
        if tag == 'fprop':
            return 'x%s k%dx%d s%d -> %s' % (tuple(a[0].shape), a[3], a[4], a[5] if len(a) > 5 else k.get('stride', 1),
                                              tuple(out.shape))
        if tag == 'wgrad':
            return 'x%s dy%s k%dx%d' % (tuple(a[0].shape), tuple(a[1].shape), a[4], a[5])
        return 'hidden%s C%d k%d' % (tuple(a[0].shape), a[9], a[10])

    def table(self, top=40):
        rows = {}
        for tag, fl, s, e, desc in self.records:
            r = rows.setdefault((tag, desc), [0.0, 0.0, 0])
            r[0] += fl; r[1] += s.elapsed_time(e); r[2] += 1
        out = []
        for (tag, desc), (fl, ms, n) in sorted(rows.items(), key=lambda kv: -kv[1][1])[:top]:
            out.append('%7.3f ms x%-3d %7.1f TF/s  %-8s %s' % (ms, n, fl / (ms * 1e-3) / 1e12 if ms > 0 else 0, tag, desc))
        return '\n'.join(out)

    def remove(self):
        from imaginaire_b200 import ops
        ops.conv2d_fprop, ops.conv2d_wgrad, ops.spade_gb_conv_modulate, ops.conv2d_fprop_window = self._orig

    def summary(self):
        by = {}
        for tag, fl, s, e, _ in self.records:
            ms = s.elapsed_time(e)
            a = by.setdefault(tag, [0.0, 0.0, 0])
            a[0] += fl; a[1] += ms; a[2] += 1
        tot_fl = sum(v[0] for v in by.values())
        tot_ms = sum(v[1] for v in by.values())
        return by, tot_fl, tot_ms


def conv_traffic_from_profile():
    """DRAM bytes of one launch of the dominant convolution shape as ncu measured them (tools/ncu_traffic.py writes
    the file from a `--set full` capture); None if the capture is not in the tree."""
    try:
        with open(os.path.join(ROOT, 'profiles', 'r02_conv_traffic.json')) as f:
            d = json.load(f)
        return float(d['dram_bytes_read']) + float(d['dram_bytes_write'])
    except (OSError, KeyError, ValueError):
        return None


def hbm_rooflines(dev, hbm_peak_gbs):
    """Second half of BASELINE.json's metric: achieved HBM GB/s of the norm / modulation / warp kernels, each timed
    alone with CUDA events (5 calls replayed as a CUDA graph after 2 warm-ups) on tensors far larger than L2: the C3 layer 8 x 512 x 128 x 256
    (bf16 NHWC) for the norm kernels, 8 x 64 x 512 x 1024 fp32 on the SURVEY 8(d) flow for the warps.  Bytes are the
    algorithmic ones of DESIGN.md section 3."""
    import torch
    from imaginaire_b200 import ops

    def timed(fn):
        # five calls recorded in a CUDA graph and replayed: the kernels run back to back, so a 50 us kernel is not
        # timed together with the host's launch gaps
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(2):
                fn()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            for _ in range(5):
                fn()
        graph.replay()
        torch.cuda.synchronize()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        graph.replay()
        e.record()
        torch.cuda.synchronize()
        return s.elapsed_time(e) / 5 * 1e-3

    out = {}

    def rec(name, nbytes, fn):
        t = timed(fn)
        out[name] = {'GBps': nbytes / t / 1e9, 'frac': nbytes / t / 1e9 / hbm_peak_gbs, 'bytes': nbytes, 'us': t * 1e6}

    g = torch.Generator(device=dev).manual_seed(7)
    n, c, h, w = 8, 512, 128, 256
    t1 = n * c * h * w * 2                      # one bf16 activation tensor
    mk = lambda: torch.randn(n, c, h, w, device=dev, generator=g).bfloat16().contiguous(memory_format=torch.channels_last)  # noqa: E731
    x, gam, bet, dy = mk(), mk(), mk(), mk()
    aff_w, aff_b = torch.ones(c, device=dev), torch.zeros(c, device=dev)
    sums = ops.bn_stats(x, False)
    mean, rstd = ops.bn_finalize(sums, n * h * w, 1e-5)
    y = ops.spade_modulate_fwd(x, mean, rstd, aff_w, aff_b, gam, bet, False, 1)
    rec('stats_kernel', t1, lambda: ops.bn_stats(x, False))
    rec('modulate_fwd (x, gamma, beta -> y)', 4 * t1, lambda: ops.spade_modulate_fwd(x, mean, rstd, aff_w, aff_b, gam, bet, False, 1))
    # backward = pass 1 (x, dy, gamma, beta -> dgamma, dbeta: 6 tensors) + pass 2 (x, dy, gamma -> dx: 4 tensors)
    rec('modulate_bwd (both passes)', 10 * t1,
        lambda: ops.spade_modulate_bwd(x, dy, mean, rstd, aff_w, aff_b, gam, bet, False, 1, n * h * w))
    rec('act_bwd_sum_kernel', 3 * t1, lambda: ops.act_bwd_bias(dy, y, 1, True))
    del x, gam, bet, dy, y
    b_, c_, hh, ww = 8, 64, 512, 1024
    img = torch.randn(b_, c_, hh, ww, device=dev, generator=g)
    flow = torch.nn.functional.interpolate(torch.randn(b_, 2, 32, 64, device=dev, generator=g) * 8.0, size=(hh, ww),
                                           mode='bilinear', align_corners=False).contiguous()
    go = torch.randn(b_, c_, hh, ww, device=dev, generator=g)
    px = b_ * hh * ww * 4
    rec('resample2d_fwd', (2 * c_ + 2) * px, lambda: ops.resample2d_fwd(img, flow))
    rec('gridsample_warp_fwd', (2 * c_ + 2) * px, lambda: ops.gridsample_warp_fwd(img, flow))
    rec('resample2d_bwd', (3 * c_ + 4) * px, lambda: ops.resample2d_bwd(img, flow, go))
    return out


# ---------------------------------------------------------------------- CPU baseline

def cpu_reference_step(cfg, batch=1, seed=0):
    """One bounded sample of the reference's CPU path: ONE REAL TRAINING ITERATION of the benchmark's configuration at
    batch 1 -- dis_update, gen_update, moving average: three generator passes, two discriminator calls on real and
    fake, hinge + feature-matching + KL + VGG-19 perceptual losses, both backward passes, both Adam steps -- in fp32 on
    all host cores through the CPU oracle port of the reference's PyTorch code (oracle/train_oracle.py, pinned to the
    reference trainer's own outputs by tests/test_train_oracle.py).  Returns (seconds, images)."""
    import torch
    from oracle import train_oracle as to
    from imaginaire_b200.discriminators.spade import Discriminator
    from imaginaire_b200.generators.spade import Generator
    from imaginaire_b200.utils.data import get_crop_h_w
    from imaginaire_b200.utils.init_weight import weights_init
    state = cpu_reference_step.__dict__
    if 'sd_g' not in state:
        from torchvision.models.vgg import vgg19
        torch.manual_seed(seed)
        net_G, net_D = Generator(cfg.gen, cfg.data), Discriminator(cfg.dis, cfg.data)
        fn = weights_init(cfg.trainer.init.type, cfg.trainer.init.gain)
        net_G.apply(fn)
        net_D.apply(fn)
        state['sd_g'], state['sd_d'] = net_G.state_dict(), net_D.state_dict()
        state['vgg'] = vgg19(weights=None).features.state_dict()     # random filters, as on the GPU arm (no download)
        state['spec'] = to.trainer_spec_from_cfg(cfg, get_crop_h_w(cfg.data.train.augmentations))
        ema = {k[:-5] if k.endswith('weight_orig') else k: None for k in state['sd_g']
               if not k.endswith(('weight_u', 'weight_v'))}
        zeros = {k: torch.zeros_like(state['sd_g'].get(k, state['sd_g'].get(k + '_orig'))) for k in ema}
        state['train'] = {'opt_d': {}, 'opt_g': {}, 'ema_updates': 0,
                          'ema': to.model_average_update(zeros, state['sd_g'], 0.0),
                          'ema_start': int(getattr(cfg.trainer, 'model_average_start_iteration', 1000)),
                          'ema_beta': float(getattr(cfg.trainer, 'model_average_beta', 0.9999))}
    data = synthetic_batch(cfg, batch, seed)
    eps_d, eps_g = torch.randn(batch, cfg.gen.style_dims), torch.randn(batch, cfg.gen.style_dims)
    t0 = time.perf_counter()
    _, _, state['sd_g'], state['sd_d'], state['train'] = to.train_iteration(
        state['sd_g'], state['sd_d'], state['vgg'], data, state['spec'], eps_d, eps_g, state['train'])
    return time.perf_counter() - t0, batch


# Algorithmic FLOPs per image (SURVEY.md section 3.1 / BASELINE.md section 3): one training
# iteration = 4 F_G + 4.5 F_D + 3 F_VGG (GFLOP).
F_G, F_D, F_VGG = 2883.0, 172.0, 102.0


def run_reference(args):
    """Reference arm: the reference's own (CPU, PyTorch) implementation of the path, as ported in
    oracle/, timed on this box's host cores.  /root/reference is Python and does not travel;
    the oracle is pinned to it by tests/golden.  Each step is a bounded sample of the workload: one whole training
    iteration of the same configuration at batch 1 (the GPU arm runs batch 8 per step)."""
    import torch
    from imaginaire_b200.config import Config
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    cfg = Config(os.path.join(ROOT, 'configs', args.config))
    steps, warmup = args.steps, max(args.warmup, 1)
    if steps + warmup > 8:   # keep the whole run within a few minutes
        warmup = 1
        steps = min(steps, 4)
    for _ in range(warmup):
        cpu_reference_step(cfg)
    t, images = 0.0, 0
    for _ in range(steps):
        dt, n = cpu_reference_step(cfg)
        t += dt
        images += n
    value = images / t
    sample = CPU_SAMPLE % (t / steps)
    line = {
        'impl': 'reference', 'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': args.gpus, 'steps': steps,
        'warmup': warmup, 'ms_per_step': 1000.0 * t / steps, 'higher_is_better': True, 'scaling': 'weak',
        'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic',
        'config': {'workload': 'SPADE G+D training step, base128 generator + SPADE discriminator, 256x512, 35 labels, '
                               'SyncBN, Adam, EMA (CPU sample: one whole training iteration at batch 1)',
                   'config_file': 'configs/' + args.config},
        'cpu_baseline': {'value': value, 'unit': UNIT, 'cores': cores, 'kind': 'port', 'sample': sample},
        'e2e': {'value': value, 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
        'gpu_launches': 0,
    }
    print(json.dumps(line), flush=True)


CPU_SAMPLE = ('one whole G+D training iteration (dis_update + gen_update + moving average, fp32) of the same '
              'configuration at batch 1, 256x512, through the CPU oracle port of the reference trainer '
              '(oracle/train_oracle.py): %.2f s per iteration')


# ------------------------------------------------------------------------- our arm

def run_ours(args):
    import torch
    import torch.distributed as dist
    from imaginaire_b200 import _C
    from imaginaire_b200.config import Config
    from imaginaire_b200.trainers.spade import Trainer
    from imaginaire_b200.utils.trainer import get_model_optimizer_and_scheduler

    world = int(os.environ.get('WORLD_SIZE', '1'))
    rank = int(os.environ.get('RANK', '0'))
    local_rank = int(os.environ.get('LOCAL_RANK', '0'))
    if not torch.cuda.is_available():
        raise RuntimeError('bench.py needs a CUDA device: the product path has no CPU fallback')
    torch.cuda.set_device(local_rank)
    dev = torch.device('cuda', local_rank)
    if world > 1:
        os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
        # collectives on a high-priority stream: their CTAs are placed before the next persistent conv kernel's
        os.environ.setdefault('TORCH_NCCL_HIGH_PRIORITY', '1')
        dist.init_process_group('nccl', init_method='env://')
    _C.lib()   # fails loudly if the CUDA library is missing

    cfg = Config(os.path.join(ROOT, 'configs', args.config))
    batch = args.batch or cfg.data.train.batch_size
    net_G, net_D, opt_G, opt_D, sch_G, sch_D = get_model_optimizer_and_scheduler(cfg, seed=0, device=dev)
    trainer = Trainer(cfg, net_G, net_D, opt_G, opt_D, sch_G, sch_D, (), None)
    net_G.train(); net_D.train()

    host = synthetic_batch(cfg, batch, seed=100 + rank, pinned=True)
    resident = {k: v.to(dev) for k, v in host.items()}
    h2d_bytes = sum(v.numel() * v.element_size() for v in host.values())

    def one_step(data):
        # Fresh tensor objects every step (no copy): per-batch caches that the layers keep on the input
        # tensors (resized / pooled label maps) are rebuilt each step, as with a new batch in training.
        return trainer.step({k: v.detach() for k, v in data.items()})

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    for _ in range(args.warmup):
        one_step(resident)
    # The iteration is captured once as a CUDA graph and replayed (utils/cuda_graph.py); every call still copies its
    # batch into the graph's input tensors.
    from torch.nn.parallel import DistributedDataParallel as _DDP
    # (with several ranks only when no DistributedDataParallel wrapper is in the loop: its autograd hooks and bucket
    # bookkeeping are host-side; the peer-memory gradient path + flat all-reduce of utils/trainer.py is all kernels)
    use_graph = not args.no_graph and os.environ.get('IMX_BENCH_GRAPH', '1') != '0' and not args.one_step and \
        not isinstance(net_G, _DDP) and not isinstance(net_D, _DDP)
    if use_graph:
        from imaginaire_b200.utils.cuda_graph import GraphedTrainerStep
        graphed = GraphedTrainerStep(trainer, resident)
        eager_step = one_step

        def one_step(data):     # noqa: F811
            return graphed(data)
        one_step(resident)      # capture + first replay, untimed
    if args.one_step:
        torch.cuda.synchronize()
        one_step(resident)
        torch.cuda.synchronize()
        print(json.dumps({'one_step': True, 'note': 'profiling run, no metric'}), flush=True)
        return

    lib = _C.lib()
    sampler = ClockSampler(local_rank)
    # Python's cyclic garbage collector pauses each rank at its own moments; with a barrier-coupled step
    # (SyncBN) every pause of any rank stalls all of them.  Collect once now, then keep the collector off
    # for the timed regions (a training loop would call gc.collect() every few hundred iterations).
    import gc
    gc.collect()
    if os.environ.get('IMX_BENCH_GC', '0') != '1':
        gc.disable()
    # ---- timed region 1: inputs resident in HBM ----
    barrier()
    if not args.no_clocks:
        sampler.start()
    lib.launches = 0
    start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    start.record()
    for _ in range(args.steps):
        one_step(resident)
    end.record()
    barrier()
    launches = lib.launches
    ms = start.elapsed_time(end)
    # ---- timed region 2: end to end (pinned host -> device copies + loss read-back each step) ----
    # The copies run on a second stream into two staging batches, so the upload of batch i + 1 overlaps the compute
    # of batch i, as a data loader with pinned memory and non_blocking copies (or the reference's prefetching
    # iterator) does it.  Every step's inputs are copied inside the timed region (the first one right after the start
    # event), and every step ends with a device -> host read of its losses.
    copy_stream = torch.cuda.Stream()
    staging = [{k: torch.empty_like(v) for k, v in resident.items()} for _ in range(2)]
    ready = [torch.cuda.Event(), torch.cuda.Event()]

    def upload(i):
        with torch.cuda.stream(copy_stream):
            for k, v in host.items():
                staging[i % 2][k].copy_(v, non_blocking=True)
            ready[i % 2].record(copy_stream)

    barrier()
    start.record()
    d2h = 0
    upload(0)
    for i in range(args.steps):
        if i + 1 < args.steps:
            upload(i + 1)       # staging[(i + 1) % 2] was last read by step i - 1, which the loss read-back has waited for
        torch.cuda.current_stream().wait_event(ready[i % 2])
        g_loss, d_loss = one_step(staging[i % 2])
        losses = torch.stack([g_loss.reshape(()), d_loss.reshape(())]).cpu()   # device -> host read of the step's result
        d2h = losses.numel() * losses.element_size()
    end.record()
    barrier()
    ms_e2e = start.elapsed_time(end)
    clocks = sampler.stop()
    if world > 1:
        t = torch.tensor([ms, ms_e2e], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, ms_e2e = float(t[0]), float(t[1])

    # ---- roofline of the dominant kernel family: one more iteration with per-launch CUDA events ----
    # An un-probed iteration is queued first and NOT waited for, so the device stays backlogged while the probed
    # one is enqueued: each start event then fires when the previous kernel ends, not when the host got there.
    if use_graph:
        one_step = eager_step       # per-launch events need the eager loop
    one_step(resident)
    probe = ConvProbe()
    probe.install()
    one_step(resident)
    torch.cuda.synchronize()
    probe.remove()
    by, conv_flops, conv_ms = probe.summary()
    if os.environ.get('IMX_BENCH_VERBOSE') and rank == 0:
        print(probe.table(int(os.environ.get("IMX_BENCH_TABLE_TOP", "40"))), file=sys.stderr, flush=True)
    peaks = load_peaks()
    achieved = conv_flops / (conv_ms * 1e-3) / 1e12 if conv_ms > 0 else 0.0
    ms_per_step = ms / args.steps
    roofline = {
        'bound': 'tensor', 'kernel': 'conv_igemm_kernel / conv_wgrad_kernel (tcgen05 implicit GEMM), all launches of one iteration',
        'achieved': achieved, 'peak': peaks['tf_sustained'], 'unit': 'TFLOP/s', 'frac': achieved / peaks['tf_sustained'],
        # dram__bytes_read.sum + dram__bytes_write.sum of the family's most frequent launch (conv_igemm_kernel<256,0,1>,
        # 3x3 512->512 at 8x128x256; algorithmic 541 MB = x 268 + filter 4.7 + y 268), read from the committed ncu
        # summary of that launch -- null when the file is not there
        'traffic': conv_traffic_from_profile(), 'traffic_unit': 'bytes per launch (ncu --set full, one launch of the dominant shape; profiles/r02_conv_traffic.json)',
        'peak_source': peaks['source'] + ' bf16_tflops_sustained (kernel timed inside a long step)',
        'conv_share_of_step': conv_ms / ms_per_step if ms_per_step > 0 else None,
        'by_kind': {k: {'tflops': v[0] / (v[1] * 1e-3) / 1e12 if v[1] > 0 else 0.0, 'ms': v[1], 'launches': v[2]}
                    for k, v in by.items()},
    }

    if os.environ.get('IMX_BENCH_CHECKSUM') and rank == 0:
        gsum = sum(float(p.detach().double().abs().sum()) for p in net_G.parameters())
        dsum = sum(float(p.detach().double().abs().sum()) for p in net_D.parameters())
        print('checksum G %.9e D %.9e' % (gsum, dsum), file=sys.stderr, flush=True)
    value = world * batch * args.steps / (ms * 1e-3)
    e2e_value = world * batch * args.steps / (ms_e2e * 1e-3)
    line = {
        'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': world, 'steps': args.steps, 'warmup': args.warmup,
        'ms_per_step': ms_per_step, 'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None,
        'dtype': 'bf16', 'data': 'synthetic',
        'config': {'workload': 'SPADE G+D training step (BASELINE configs[2]): base128 generator + SPADE discriminator, '
                               '256x512, 35-class one-hot labels, batch %d/GPU, SyncBN, Adam, EMA' % batch,
                   'config_file': 'configs/' + args.config, 'global_batch': world * batch,
                   'parallelism': 'dp%d' % world, 'cuda_graph': bool(use_graph), 'l2': 'working set (GBs of activations per step) is far larger than the 126 MB L2; no explicit flush'},
        'e2e': {'value': e2e_value, 'unit': UNIT, 'h2d_bytes_per_step': h2d_bytes, 'd2h_bytes_per_step': d2h,
                'ms_per_step': ms_e2e / args.steps},
        'gpu_launches': launches,
        'roofline': roofline,
        'clocks': clocks,
        'roofline_hbm': hbm_rooflines(dev, peaks['hbm']) if rank == 0 and not args.no_hbm_probe else None,
        'step_tflops': world * batch * (4 * F_G + 4.5 * F_D + 3 * F_VGG) / 1e3 / (ms_per_step * 1e-3) / max(world, 1),
    }
    if rank == 0:
        if world == 1 and not args.no_cpu_baseline:
            cores = os.cpu_count() or 1
            torch.set_num_threads(cores)
            cpu_reference_step(cfg)            # warm-up (weight build, thread pool, first Adam step)
            dt, n = cpu_reference_step(cfg)
            line['cpu_baseline'] = {'value': n / dt, 'unit': UNIT, 'cores': cores, 'kind': 'port', 'sample': CPU_SAMPLE % dt}
        print(json.dumps(line), flush=True)
    if use_graph:
        graphed.close()
    if world > 1:
        # a teardown that hangs (a communicator waiting for a peer that already left) must not hold the GPUs: the result
        # line is out, so after a grace period the process just ends
        import threading
        t = threading.Timer(45.0, lambda: os._exit(0))
        t.daemon = True
        t.start()
        dist.barrier()
        dist.destroy_process_group()


if __name__ == '__main__':
    a = parse_args()
    if a.impl == 'reference':
        run_reference(a)
    else:
        run_ours(a)
