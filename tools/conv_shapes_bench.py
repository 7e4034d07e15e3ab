"""Replays every distinct conv launch of one training iteration back to back (developer tool).

One training iteration is run with recording wrappers around the three conv entry points; every
distinct (entry point, shapes, flags) signature is then replayed REPS times back to back between
two CUDA events, so the per-shape numbers carry no host gaps.  Prints count x time, sorted by the
time the shape costs per training iteration.
"""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from imaginaire_b200 import ops
from imaginaire_b200.config import Config
from imaginaire_b200.trainers.spade import Trainer
from imaginaire_b200.utils.trainer import get_model_optimizer_and_scheduler

REPS = 10


def clone_args(a, k):
    def c(v):
        if torch.is_tensor(v):
            t = torch.empty_like(v)
            if v.dtype.is_floating_point:
                t.normal_()
            else:
                t.zero_()
            return t
        return v
    return tuple(c(v) for v in a), {kk: c(v) for kk, v in k.items()}


def sig(a, k):
    def s(v):
        if torch.is_tensor(v):
            return ('T', tuple(v.shape), str(v.dtype))
        return v
    return tuple(s(v) for v in a) + tuple((kk, s(v)) for kk, v in sorted(k.items()))


def main():
    batch = int(sys.argv[1]) if len(sys.argv) > 1 else 8
    cfg = Config(os.path.join(bench.ROOT, 'configs', bench.CONFIG_NAME))
    dev = torch.device('cuda:0')
    net_G, net_D, opt_G, opt_D, sch_G, sch_D = get_model_optimizer_and_scheduler(cfg, seed=0, device=dev)
    trainer = Trainer(cfg, net_G, net_D, opt_G, opt_D, sch_G, sch_D, (), None)
    data = {k: v.to(dev) for k, v in bench.synthetic_batch(cfg, batch, 100).items()}
    trainer.step(dict(data))
    seen = {}
    orig = {n: getattr(ops, n) for n in ('conv2d_fprop', 'conv2d_wgrad', 'spade_gb_conv_modulate')}
    probe = bench.ConvProbe()

    def rec(name):
        fn = orig[name]

        def w(*a, **k):
            key = (name, sig(a, k))
            if key not in seen:
                seen[key] = [0, a, k]
            seen[key][0] += 1
            return fn(*a, **k)
        return w
    for n in orig:
        setattr(ops, n, rec(n))
    trainer.step(dict(data))
    for n, f in orig.items():
        setattr(ops, n, f)
    torch.cuda.synchronize()
    shapes = [(key, cnt, a, k) for key, (cnt, a, k) in seen.items()]
    del trainer, net_G, net_D, opt_G, opt_D
    torch.cuda.empty_cache()
    rows = []
    for (name, _), cnt, a, k in shapes:
        a2, k2 = clone_args(a, k)
        fn = orig[name]
        out = fn(*a2, **k2)
        if name == 'conv2d_fprop':
            n, _, ho, wo = out.shape
            fl = 2.0 * n * ho * wo * a2[2] * a2[0].shape[1] * a2[3] * a2[4]
            desc = bench.ConvProbe.describe('fprop', out, a2, k2) + ' flags=%s' % k2.get('flags', 0)
            if k2.get('stats') is not None:
                desc += ' +stats'
            if k2.get('residual') is not None:
                desc += ' +res'
        elif name == 'conv2d_wgrad':
            n, _, ho, wo = a2[1].shape
            fl = 2.0 * n * ho * wo * a2[3] * a2[2] * a2[4] * a2[5]
            desc = bench.ConvProbe.describe('wgrad', out, a2, k2)
        else:
            n, chp, h, wd = a2[0].shape
            fl = 2.0 * n * h * wd * (2 * a2[9]) * chp * a2[10] * a2[11]
            desc = bench.ConvProbe.describe('spade_gb', out, a2, k2)
        del out
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        s.record()
        for _ in range(REPS):
            fn(*a2, **k2)
        e.record()
        torch.cuda.synchronize()
        ms = s.elapsed_time(e) / REPS
        rows.append((ms * cnt, ms, cnt, fl, name, desc))
        del a2, k2
    rows.sort(reverse=True)
    tot_ms = sum(r[0] for r in rows)
    tot_fl = sum(r[3] * r[2] for r in rows)
    print('conv time per iteration %.2f ms, %.1f TFLOP, %.1f TF/s aggregate, %d distinct shapes' %
          (tot_ms, tot_fl / 1e12, tot_fl / tot_ms / 1e9, len(rows)))
    ideal = 1392.7e12
    print('%8s %8s %4s %8s %8s  %s' % ('tot ms', 'ms', 'x', 'TF/s', 'lost ms', 'shape'))
    for tot, ms, cnt, fl, name, desc in rows[:int(os.environ.get('TOP', 70))]:
        lost = tot - cnt * fl / ideal * 1e3
        print('%8.3f %8.4f %4d %8.1f %8.3f  %-9s %s' % (tot, ms, cnt, fl / ms / 1e9, lost, name.replace('conv2d_', ''), desc))


if __name__ == '__main__':
    main()
