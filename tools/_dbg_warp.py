import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, numpy as np
from imaginaire_b200 import ops
from tests import helpers as H
from oracle import warp_oracle as wo
b,c,h,w = 1,16,512,1024
g = torch.Generator().manual_seed(21)
img = torch.randn(b,c,h,w,generator=g); flow = H.synthetic_flow(22,b,h,w,scale=float(sys.argv[1]) if len(sys.argv)>1 else 8.0)
import time
ic, fc = img.cuda(), flow.cuda(); torch.cuda.synchronize(); t0 = time.time()
try:
    out = ops.resample2d_fwd(ic, fc)
    torch.cuda.synchronize()
finally:
    print('kernel wall %.3f s' % (time.time() - t0), flush=True)
ref = wo.resample2d_forward(img.numpy(), flow.numpy())
print('equal', np.array_equal(out.cpu().numpy(), ref), np.abs(out.cpu().numpy()-ref).max())
