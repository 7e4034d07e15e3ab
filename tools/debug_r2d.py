import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from oracle import warp_oracle as wo
from tests import helpers as H
from imaginaire_b200 import ops
b, c, h, w = 2, 3, 64, 128
g = torch.Generator().manual_seed(11)
img = torch.randn(b, c, h, w, generator=g)
flow = H.synthetic_flow(12, b, h, w)
out = ops.resample2d_fwd(img.cuda(), flow.cuda()).cpu().numpy()
ref = wo.resample2d_forward(img.numpy(), flow.numpy())
bad = np.argwhere(out != ref)
print('mismatches', len(bad), 'of', out.size, 'maxabs', np.abs(out - ref).max())
for idx in bad[:8]:
    bb, cc, y, x = idx
    dx, dy = flow[bb, 0, y, x].item(), flow[bb, 1, y, x].item()
    xf = np.float32(x) + np.float32(dx); yf = np.float32(y) + np.float32(dy)
    print(idx, 'flow', dx, dy, 'alpha', xf - np.floor(xf), 'beta', yf - np.floor(yf), 'gpu', out[bb, cc, y, x], 'cpu', ref[bb, cc, y, x],
          'diff', out[bb, cc, y, x] - ref[bb, cc, y, x])
rows = np.unique(bad[:, 2]); print('rows with mismatches', rows[:40])
