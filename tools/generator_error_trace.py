"""Per-block error of the native SPADE generator against the fp32 oracle, beside what bf16 itself costs.

    python tools/generator_error_trace.py [--out profiles/r02_generator_error_trace.md]     (on the GPU box)

For configs/spade_unit_test.yaml (BASELINE configs[0], joint projection, instance norm) and configs/spade_c3_mini.yaml
(the C3 architecture: sync batch norm, separate 5x5 projections, spectral norm everywhere) it runs the same seeded
network and inputs through
  * the fp32 CPU oracle (oracle/spade_oracle.py, pinned to the reference by tests/golden/generator_*.npz),
  * the oracle with only the convolution OPERANDS rounded to bf16 ("operands"),
  * the oracle with every tensor the native pipeline stores between kernels rounded to bf16 as well ("storage": one
    rounding when a convolution output is written, one after the modulation -- the statistics of a batch / instance
    norm are not known when its input is produced, so a bf16-storage pipeline cannot avoid the first one),
  * the CUDA path, no-grad (gamma / beta stay in fp32 in tensor memory) and grad (gamma / beta stored in bf16),
and prints, block by block, the relative L2 error of each against fp32.  The oracle is the checker here, never part
of the product path.
"""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import spade_oracle as so  # noqa: E402
from tests import helpers as H  # noqa: E402


def rel(a, b):
    return float((a.double() - b.double()).norm() / (b.double().norm() + 1e-30))


def oracle_run(sd, data, spec, eps, quant, store, store_gb=True):
    bf = (lambda t: t.bfloat16().float())
    so.QUANT, so.STORE, so.STORE_GAMMA_BETA, so.TRACE = (bf if quant else None), (bf if store else None), store_gb, []
    try:
        out = so.generator(sd, data, spec, True, eps)
        return dict(so.TRACE), out['fake_images']
    finally:
        so.QUANT, so.STORE, so.STORE_GAMMA_BETA, so.TRACE = None, None, True, None


def gpu_run(build, data, eps, grad):
    from imaginaire_b200 import functional as NF
    from imaginaire_b200.generators import spade as gen
    dev = torch.device('cuda:0')
    net = build().to(dev).train()
    net.style_encoder.freeze_random = True
    net.style_encoder.eps = eps.to(dev)
    gen.TRACE = []
    try:
        with torch.set_grad_enabled(grad):
            out = net({k: v.to(dev) for k, v in data.items()})
        trace = {}
        for name, t in gen.TRACE:
            c = getattr(net.spade_generator, name)._out_channels if hasattr(getattr(net.spade_generator, name), '_out_channels') else t.shape[1]
            trace[name] = NF.to_nchw(t.detach(), t.shape[1]).float().cpu()
        return trace, out['fake_images'].detach().float().cpu()
    finally:
        gen.TRACE = None


def run_config(name, config, crop, data, eps, build, lines):
    from imaginaire_b200.config import Config
    cfg = Config(os.path.join(ROOT, 'configs', config))
    sd = build().state_dict()
    spec = so.generator_spec_from_cfg(cfg.gen, crop)
    ref_t, ref = oracle_run(sd, data, spec, eps, False, False)
    runs = [('operands', *oracle_run(sd, data, spec, eps, True, False)),
            ('storage (no-grad schedule)', *oracle_run(sd, data, spec, eps, True, True, store_gb=False)),
            ('storage (grad schedule)', *oracle_run(sd, data, spec, eps, True, True, store_gb=True)),
            ('CUDA no-grad', *gpu_run(build, data, eps, False)),
            ('CUDA grad', *gpu_run(build, data, eps, True))]
    lines.append('\n## %s (%s)\n' % (name, config))
    lines.append('Relative L2 error against the fp32 oracle after every block; last row: the image after tanh '
                 '(max-abs / mean-abs).\n')
    lines.append('| block | ' + ' | '.join(r[0] for r in runs) + ' | CUDA no-grad / storage | CUDA grad / storage |')
    lines.append('|---|' + '---|' * (len(runs) + 2))
    last = None
    for blk, ref_x in ref_t.items():
        row, vals = [], {}
        for tag, tr, _ in runs:
            x = tr.get(blk)
            if x is None:
                row.append('-')
                continue
            x = x[:, :ref_x.shape[1]]
            if x.shape != ref_x.shape:
                row.append('shape?')
                continue
            # the last residual block hands the to-RGB block's LeakyReLU its output already activated (native path)
            e = min(rel(x, ref_x), rel(x, so.leaky_relu(ref_x))) if tag.startswith('CUDA') else rel(x, ref_x)
            vals[tag] = e
            row.append('%.4f' % e)
        r1 = vals.get('CUDA no-grad', 0) / max(vals.get('storage (no-grad schedule)', 1e-30), 1e-30)
        r2 = vals.get('CUDA grad', 0) / max(vals.get('storage (grad schedule)', 1e-30), 1e-30)
        lines.append('| %s | %s | %.2f | %.2f |' % (blk, ' | '.join(row), r1, r2))
        last = (r1, r2)
    row = []
    for tag, _, img in runs:
        d = (img - ref).abs()
        row.append('%.4f / %.5f' % (float(d.max()), float(d.mean())))
    lines.append('| image (max-abs / mean-abs) | ' + ' | '.join(row) + ' | | |')
    return last


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--out', default=None)
    args = ap.parse_args()
    lines = ['# Generator error trace: CUDA (bf16 storage, fp32 accumulate) vs the fp32 oracle, block by block',
             '', 'Produced by `tools/generator_error_trace.py` on a B200. "operands": fp32 oracle with only the '
             'convolution operands rounded to bf16; "storage": also every tensor the pipeline keeps in HBM between '
             'kernels (see the tool\'s docstring). The ratio columns say how much of the CUDA path\'s error bf16 '
             'storage itself explains (1.0 = all of it).']
    d = H.load_golden('generator_unit_test.npz')
    label, images = H.unit_test_generator_inputs()
    run_config('unit-test generator', 'spade_unit_test.yaml', (256, 256), {'label': label, 'images': images},
               torch.from_numpy(d['eps']), lambda: H.build_generator('spade_unit_test.yaml')[1], lines)

    d = H.load_golden('generator_c3mini.npz')

    def build_c3():
        from imaginaire_b200.config import Config
        from imaginaire_b200.discriminators.spade import Discriminator
        from imaginaire_b200.generators.spade import Generator
        from imaginaire_b200.utils.init_weight import weights_init
        cfg = Config(os.path.join(ROOT, 'configs', 'spade_c3_mini.yaml'))
        torch.manual_seed(0)
        net = Generator(cfg.gen, cfg.data)
        Discriminator(cfg.dis, cfg.data)
        net.apply(weights_init(cfg.trainer.init.type, cfg.trainer.init.gain))
        return net
    run_config('C3-style generator, 16 filters', 'spade_c3_mini.yaml', (256, 256), H.c3mini_batch(),
               torch.from_numpy(d['eps']), build_c3, lines)
    text = '\n'.join(lines) + '\n'
    print(text)
    if args.out:
        with open(args.out, 'w') as f:
            f.write(text)


if __name__ == '__main__':
    main()
