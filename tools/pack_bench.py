"""Bandwidth of the filter pack kernels on the C3 generator's largest filters (developer tool; one GPU)."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from imaginaire_b200 import ops

PEAK = 6549.4
try:
    with open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'MEASURED_PEAKS.json')) as f:
        PEAK = json.load(f)['hbm_gbs']
except OSError:
    pass


def timeit(fn, iters=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters):
        fn()
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) / iters * 1e-3


SHAPES = [(2048, 2048, 3), (1024, 1024, 3), (2048, 128, 5), (128, 2048, 5), (512, 512, 3), (512, 256, 4), (128, 40, 5)]
if '--one' in sys.argv:      # for ncu: the largest filter only
    SHAPES = SHAPES[:1]
for cout, cin, k in SHAPES:
    w = torch.randn(cout, cin, k, k, device='cuda')
    ws = [torch.randn(cout, cin, k, k, device='cuda') for _ in range(max(1, int(3e8 // (w.numel() * 4))))]   # > L2 in rotation
    nbytes = w.numel() * 6
    for flip in (False, True):
        it = iter(range(10 ** 9))
        t = timeit(lambda: ops.pack_conv_weight(ws[next(it) % len(ws)], transpose_flip=flip))
        print('pack %4d x %4d x %dx%d %-5s %8.1f us  %6.0f GB/s  %.2f of the HBM peak' %
              (cout, cin, k, k, 'flip' if flip else 'plain', t * 1e6, nbytes / t / 1e9, nbytes / t / 1e9 / PEAK), flush=True)
wg, wb = torch.randn(2048, 128, 5, 5, device='cuda'), torch.randn(2048, 128, 5, 5, device='cuda')
t = timeit(lambda: ops.pack_spade_gb_weight(wg, wb))
print('interleaved gamma/beta pack 2 x 2048 x 128 x 5x5 %8.1f us  %6.0f GB/s' % (t * 1e6, 2 * wg.numel() * 6 / t / 1e9))
