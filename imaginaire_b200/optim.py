"""Optimizers of the training step on the library's own kernels.

ref: get_optimizer_for_params, imaginaire/utils/trainer.py:261-306 ('adam' -> apex FusedAdam when
``fused_opt`` else torch.optim.Adam).  apex is not part of this build; FusedAdam here is
torch.optim.Adam's state layout and hyper-parameters with ``step()`` running
imx_adam_multi_tensor_f32 (csrc/optim.cu): one 28-byte-per-parameter pass, ~12 launches for the
generator's ~480 tensors.
"""
import ctypes

import torch

from . import _C, ops


class FusedAdam(torch.optim.Adam):
    """Drop-in for torch.optim.Adam (no amsgrad / maximize / capturable): same ``state_dict()``
    entries (``step`` as a CPU scalar tensor, ``exp_avg``, ``exp_avg_sq``), so checkpoints written
    by either load into the other."""

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0, capturable=False):
        super().__init__(params, lr=lr, betas=betas, eps=eps, weight_decay=weight_decay)
        # capturable: the step counters live on the device (one fp32 per parameter, torch.optim.Adam(capturable=True)'s
        # layout) and the kernel forms the bias corrections itself, so a step captured in a CUDA graph stays right as
        # the counters advance.  make_capturable() converts an optimizer that has already stepped.
        self.imx_capturable = capturable

    def make_capturable(self):
        self.imx_capturable = True
        for group in self.param_groups:
            for p in group['params']:
                st = self.state.get(p)
                if st and not st['step'].is_cuda:
                    st['step'] = st['step'].to(device=p.device, dtype=torch.float32)
        return self

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        for group in self.param_groups:
            by_step, dev_items = {}, []
            for p in group['params']:
                g = p.grad
                if g is None:
                    continue
                if g.is_sparse:
                    raise RuntimeError('FusedAdam does not support sparse gradients')
                if not p.is_cuda or p.dtype != torch.float32 or not p.is_contiguous():
                    raise RuntimeError('FusedAdam needs contiguous float32 CUDA parameters (no CPU fallback); got %s %s'
                                       % (p.device.type, p.dtype))
                state = self.state[p]
                if len(state) == 0:
                    state['step'] = torch.zeros((), dtype=torch.float32, device=p.device if self.imx_capturable else 'cpu')
                    state['exp_avg'] = torch.zeros_like(p, memory_format=torch.contiguous_format)
                    state['exp_avg_sq'] = torch.zeros_like(p, memory_format=torch.contiguous_format)
                if g.dtype != torch.float32 or not g.is_contiguous():
                    g = g.float().contiguous()
                if state['step'].is_cuda:
                    dev_items.append((p, g, state['exp_avg'], state['exp_avg_sq'], state['step']))
                else:
                    state['step'] += 1
                    by_step.setdefault(int(state['step']), []).append((p, g, state['exp_avg'], state['exp_avg_sq']))
            beta1, beta2 = group['betas']
            hyper = (float(group['lr']), float(beta1), float(beta2), float(group['eps']), float(group['weight_decay']))
            for step, items in by_step.items():
                _adam_launch(items, *hyper, step)
            if dev_items:
                torch._foreach_add_([t[4] for t in dev_items], 1.0)
                _adam_launch(dev_items, *hyper, None)
        return loss


def _adam_launch(items, lr, beta1, beta2, eps, weight_decay, step):
    n = len(items)
    ptr_t = ctypes.c_void_p * n
    cols = [ptr_t(*[t[k].data_ptr() for t in items]) for k in range(len(items[0]))]
    numel = (ctypes.c_int64 * n)(*[t[0].numel() for t in items])
    lib = _C.lib()
    if step is None:     # device-side step counters (fifth column)
        lib.call('imx_adam_multi_tensor_f32_capturable', cols[0], cols[1], cols[2], cols[3], cols[4], numel, n, lr, beta1,
                 beta2, eps, weight_decay, ops._stream())
    else:
        lib.call('imx_adam_multi_tensor_f32', cols[0], cols[1], cols[2], cols[3], numel, n, lr, beta1, beta2, eps,
                 weight_decay, step, ops._stream())
    # launches issued: one per 96 tensors / 1184 chunks of 32768 elements
    chunks = sum((t[0].numel() + 32767) // 32768 for t in items)
    lib.launches += max((n + 95) // 96, (chunks + 1183) // 1184) - 1
    # The kernel writes through raw pointers, which autograd's version counters do not see.  Everything that
    # caches a value derived from a parameter (packed bf16 filters in functional._packed_weight, the
    # identity-affine check in layers/activation_norm.py) keys on Tensor._version, so the counters are
    # advanced here exactly as an in-place torch update would advance them.
    torch._C._increment_version([t[0] for t in items])
    for t in items:
        t[0]._imx_trained = True      # see layers/activation_norm.py:_frozen_identity_affine
