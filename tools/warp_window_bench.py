"""Forward flow warps with and without the TMA-staged shared-memory window (developer tool; one GPU).

    FLOWGRID=32,64 IMX_WARPWIN=0|1 IMX_WARPWIN_R2D=0|1 IMX_WARPWIN_CPB=32 python tools/warp_window_bench.py

FLOWGRID: size of the random grid the flow field is interpolated from (smaller = smoother flow); the table in
profiles/r02_warp_window.md is a sweep over it.
"""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from imaginaire_b200 import ops
torch.manual_seed(0)
B, C = 8, 64
img = torch.randn(B, C, 512, 1024, device='cuda')
gh, gw = [int(v) for v in os.environ.get('FLOWGRID', '32,64').split(',')]
flow = torch.nn.functional.interpolate(torch.randn(B, 2, gh, gw, device='cuda') * 8.0, size=(512, 1024), mode='bilinear', align_corners=False).contiguous()
def t(fn, reps=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(reps): fn()
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) / reps * 1e3
bytes_ = (2 * C + 2) * B * 512 * 1024 * 4
for name, fn in (('resample2d', lambda: ops.resample2d_fwd(img, flow)), ('gridsample', lambda: ops.gridsample_warp_fwd(img, flow))):
    us = t(fn)
    print('grid', os.environ.get('FLOWGRID'), 'win', os.environ.get('IMX_WARPWIN'), 'cpb', os.environ.get('IMX_WARPWIN_CPB'), name, '%.1f us  %.3f of 6544 GB/s' % (us, bytes_ / us / 1e3 / 6544))
