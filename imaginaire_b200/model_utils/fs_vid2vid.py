"""The flow warp the vid2vid generators call, on the native kernel.

ref: imaginaire/model_utils/fs_vid2vid.py:14-39 (resample), :42-80 (get_grid)

``resample(image, flow)`` returns exactly what the reference's
F.grid_sample(image, get_grid + flow / ((size-1)/2), bilinear, border, align_corners=True)
returns (bit-exact against torch's CPU result for fp32 inputs), but the sampling grid is
never built: one kernel reads the flow, forms the coordinates with the same fp32
operation order and gathers.  ``get_grid`` is kept for callers that want the grid; it
builds it on the requested device instead of ``.to('cuda')``.
"""
import torch
from torch.autograd import Function

from .. import ops


class _GridSampleWarp(Function):
    @staticmethod
    def forward(ctx, image, flow):
        ctx.save_for_backward(image, flow)
        return ops.gridsample_warp_fwd(image, flow)

    @staticmethod
    def backward(ctx, grad_output):
        image, flow = ctx.saved_tensors
        gi, gf = ops.gridsample_warp_bwd(image, flow, grad_output.contiguous(),
                                         need_img=ctx.needs_input_grad[0], need_flow=ctx.needs_input_grad[1])
        return gi, gf


def resample(image, flow):
    """Warps ``image`` (N x C x H x W) by ``flow`` (N x 2 x H x W, pixels; channel 0 = x)."""
    assert flow.shape[1] == 2
    in_dtype = image.dtype
    out = _GridSampleWarp.apply(image.float().contiguous(), flow.float().contiguous())
    return out.to(in_dtype)


def get_grid(batchsize, size, minval=-1.0, maxval=1.0, device='cuda'):
    if len(size) == 2:
        rows, cols = size
    elif len(size) == 3:
        deps, rows, cols = size
    else:
        raise ValueError('Dimension can only be 2 or 3.')
    x = torch.linspace(minval, maxval, cols).view(1, 1, 1, cols).expand(batchsize, 1, rows, cols)
    y = torch.linspace(minval, maxval, rows).view(1, 1, rows, 1).expand(batchsize, 1, rows, cols)
    grid = torch.cat([x, y], dim=1)
    if len(size) == 3:
        z = torch.linspace(minval, maxval, deps).view(1, 1, deps, 1, 1).expand(batchsize, 1, deps, rows, cols)
        grid = torch.cat([grid.unsqueeze(2).expand(batchsize, 2, deps, rows, cols), z], dim=1)
    grid.requires_grad = False
    return grid.to(device)
