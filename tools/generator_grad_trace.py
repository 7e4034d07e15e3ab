"""Per-parameter gradient agreement of the native SPADE generator with the fp32 oracle, beside what bf16 storage
itself costs in the backward pass.

    python tools/generator_grad_trace.py [--out profiles/r02_generator_grad_trace.md]      (on the GPU box)

The C3-style generator of configs/spade_c3_mini.yaml, seeded as in tests/golden/generator_c3mini.npz, is run forward and
backward on a fixed linear loss  L = <fake_images, R> / numel  three ways:
  * fp32 CPU oracle under torch autograd (oracle/spade_oracle.py),
  * the same with every tensor the native pipeline stores in HBM rounded to bf16 in the forward pass AND its
    gradient rounded to bf16 in the backward pass (a straight-through rounding function at the same points),
  * the CUDA path.
For every parameter it prints the cosine between each of the last two and the fp32 gradient.  Parameters whose true
gradient is zero (a bias that a batch norm removes again) are listed apart.  The oracle is the checker, never part of
the product path.
"""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import spade_oracle as so  # noqa: E402
from tests import helpers as H  # noqa: E402


class RoundBoth(torch.autograd.Function):
    """bf16 rounding of a stored tensor; its gradient is stored in bf16 as well."""

    @staticmethod
    def forward(ctx, t):
        return t.bfloat16().float()

    @staticmethod
    def backward(ctx, g):
        return g.bfloat16().float()


class RoundForward(torch.autograd.Function):
    """bf16 rounding of a convolution operand; the operand's gradient is whatever the consumer stores."""

    @staticmethod
    def forward(ctx, t):
        return t.bfloat16().float()

    @staticmethod
    def backward(ctx, g):
        return g


def cos(a, b):
    a, b = a.double().reshape(-1), b.double().reshape(-1)
    return float((a * b).sum() / (a.norm() * b.norm() + 1e-300))


def build():
    from imaginaire_b200.config import Config
    from imaginaire_b200.discriminators.spade import Discriminator
    from imaginaire_b200.generators.spade import Generator
    from imaginaire_b200.utils.init_weight import weights_init
    cfg = Config(os.path.join(ROOT, 'configs', 'spade_c3_mini.yaml'))
    torch.manual_seed(0)
    net = Generator(cfg.gen, cfg.data)
    Discriminator(cfg.dis, cfg.data)
    net.apply(weights_init(cfg.trainer.init.type, cfg.trainer.init.gain))
    return cfg, net


def oracle_grads(cfg, sd, data, eps, R, emulate):
    leaves = {k: v.clone().requires_grad_(v.is_floating_point() and not k.endswith(('weight_u', 'weight_v', 'running_mean',
                                                                                     'running_var')))
              for k, v in sd.items()}
    so.QUANT, so.STORE = (RoundForward.apply, RoundBoth.apply) if emulate else (None, None)
    try:
        spec = so.generator_spec_from_cfg(cfg.gen, (256, 256))
        mu = logvar = z = None
        work = dict(leaves)
        mu, logvar, z = so.style_encoder(work, 'style_encoder.', data['images'], spec['style_kernel_size'], True, eps)
        img = so.spade_generator(work, data['label'], z, spec, True)
        loss = (img * R).sum() / R.numel()
        loss.backward()
    finally:
        so.QUANT, so.STORE = None, None
    return {k: v.grad for k, v in leaves.items() if v.grad is not None}, float(loss)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--out', default=None)
    args = ap.parse_args()
    d = H.load_golden('generator_c3mini.npz')
    eps = torch.from_numpy(d['eps'])
    data = H.c3mini_batch()
    g = torch.Generator().manual_seed(99)
    R = torch.randn(2, 3, 256, 256, generator=g)
    cfg, net = build()
    sd = {k: v.clone() for k, v in net.state_dict().items()}
    ref, loss_ref = oracle_grads(cfg, sd, data, eps, R, False)
    emu, loss_emu = oracle_grads(cfg, sd, data, eps, R, True)

    dev = torch.device('cuda:0')
    net = net.to(dev).train()
    net.style_encoder.freeze_random = True
    net.style_encoder.eps = eps.to(dev)
    for p in net.parameters():
        p.requires_grad_(True)
    out = net({k: v.to(dev) for k, v in data.items()})
    loss = (out['fake_images'].float() * R.to(dev)).sum() / R.numel()
    loss.backward()
    gpu = {k: p.grad.float().cpu() for k, p in net.named_parameters() if p.grad is not None}

    lines = ['# Generator gradients: CUDA (bf16 storage) vs the fp32 oracle, parameter by parameter', '',
             'Produced by `tools/generator_grad_trace.py` on a B200: C3-style generator (configs/spade_c3_mini.yaml), '
             'loss = <image, R> / numel. "bf16 emulation" = fp32 oracle with every HBM-resident tensor and its '
             'gradient rounded to bf16. Loss: fp32 %.6f, emulation %.6f, CUDA %.6f.' % (loss_ref, loss_emu, float(loss)),
             '', '| parameter | |g| fp32 | cos(bf16 emulation, fp32) | cos(CUDA, fp32) | |g_CUDA| / |g_fp32| |', '|---|---|---|---|---|']
    scale = max(float(v.norm()) for v in ref.values())
    zero = []
    worst_gap = 0.0
    for k, gr in ref.items():
        if k not in gpu:
            lines.append('| %s | %.3e | - | MISSING | - |' % (k, float(gr.norm())))
            continue
        if float(gr.norm()) < 1e-5 * scale:
            zero.append((k, float(gr.norm()), float(gpu[k].norm())))
            continue
        ce, cg = cos(emu[k], gr), cos(gpu[k], gr)
        worst_gap = max(worst_gap, ce - cg)
        lines.append('| %s | %.3e | %.4f | %.4f | %.3f |' % (k, float(gr.norm()), ce, cg, float(gpu[k].norm() / gr.norm())))
    lines += ['', 'Largest shortfall of the CUDA cosine below the emulation\'s: %.4f.' % worst_gap, '',
              'Parameters whose true gradient is zero (fp32 norm below 1e-5 of the largest; both columns are rounding '
              'residue): ' + ', '.join('%s (fp32 %.1e, CUDA %.1e)' % z for z in zero)]
    text = '\n'.join(lines) + '\n'
    print(text)
    if args.out:
        with open(args.out, 'w') as f:
            f.write(text)


if __name__ == '__main__':
    main()
