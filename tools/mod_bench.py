"""Times the statistics, activation-backward and the three SPADE modulate kernels on the shapes
of the C3 training step (developer tool).  Prints algorithmic GB/s and the fraction of the HBM peak."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from imaginaire_b200 import _C, ops

HBM = bench.load_peaks()['hbm']


def nhwc(n, c, h, w):
    return torch.randn(n, c, h, w, device='cuda').bfloat16().contiguous(memory_format=torch.channels_last)


def timed(fn, reps=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(reps):
        fn()
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) / reps * 1e-3


def main():
    torch.manual_seed(0)
    p = ops._ptr
    for n, c, h, w in ((8, 64, 256, 512), (8, 128, 256, 512), (8, 256, 128, 256), (8, 512, 128, 256), (8, 1024, 32, 64),
                       (8, 1024, 16, 32)):
        x, g, b, dy = (nhwc(n, c, h, w) for _ in range(4))
        sums = ops.bn_stats(x, False)
        mean, rstd = ops.bn_finalize(sums, n * h * w, 1e-5)
        y, dg, db, dx = (ops.empty_nhwc(n, c, h, w, x.device) for _ in range(4))
        red = torch.zeros(2, 1, c, device='cuda')
        gbs = torch.zeros(2, c, device='cuda')
        lib = _C.lib()
        st = ops._stream()
        elems = n * c * h * w * 2
        bg = torch.zeros(c, device='cuda')
        runs = (
            ('stats', elems, lambda: lib.call('imx_bn_stats_nhwc_bf16', p(x), p(red[0]), p(red[1]), n, h * w, c, 0, st)),
            ('act_bwd', 3 * elems, lambda: lib.call('imx_act_bwd_bias_nhwc_bf16', p(dy), p(x), p(dx), p(bg), n * h * w, c, 1, st)),
            ('fwd', 4 * elems, lambda: lib.call('imx_spade_modulate_fwd', p(x), p(mean), p(rstd), None, None, p(g), p(b), p(y),
                                                n, h, w, c, 0, 0, 1, 0, st)),
            ('bwd_reduce', 6 * elems, lambda: lib.call('imx_spade_modulate_bwd_reduce', p(x), p(dy), p(mean), p(rstd), None, None,
                                                       p(g), p(b), p(dg), p(db), p(red), p(gbs), n, h, w, c, 0, 0, 1, st)),
            ('bwd_dx', 5 * elems, lambda: lib.call('imx_spade_modulate_bwd_dx', p(x), p(dy), p(mean), p(rstd), None, None,
                                                   p(g), p(b), p(red), float(n * h * w), p(dx), n, h, w, c, 0, 0, 1, 1, st)),
        )
        for name, nbytes, fn in runs:
            t = timed(fn)
            print('%-10s N=%d C=%4d %3dx%3d  %8.1f us  %7.0f GB/s  %.2f of HBM peak' %
                  (name, n, c, h, w, t * 1e6, nbytes / t / 1e9, nbytes / t / 1e9 / HBM))
        del x, g, b, dy, y, dg, db, dx


if __name__ == '__main__':
    main()
