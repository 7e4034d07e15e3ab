"""Exponential moving average of the generator.  ref: imaginaire/utils/model_average.py:35-197

Same wrapper shape (``.module`` / ``.averaged_model`` / ``num_updates_tracked``, so the
checkpoint prefixes match) and the same update rule, including ``remove_sn``: the averaged
copy stores W / sigma under ``weight`` instead of (weight_orig, weight_u, weight_v).  The
update itself is batched with torch._foreach ops instead of one small kernel per tensor.
"""
import copy

import torch
from torch import nn
from torch.nn.utils import remove_spectral_norm


def requires_grad(model, flag=True):
    for p in model.parameters():
        p.requires_grad = flag


class ModelAverage(nn.Module):
    def __init__(self, module, beta=0.9999, start_iteration=1000, remove_sn=True):
        super().__init__()
        self.module = module
        self.averaged_model = copy.deepcopy(self.module)
        self.beta = beta
        self.remove_sn = remove_sn
        self.start_iteration = start_iteration
        self.register_buffer('num_updates_tracked', torch.tensor(0, dtype=torch.long))
        # host mirror of num_updates_tracked (the checkpointed counter, model_average.py:87-91); None = not read
        # yet / a state dict was loaded: the next update reads the buffer once (.item()), later ones stay
        # sync-free
        self._steps = None
        requires_grad(self.averaged_model, False)
        if remove_sn:
            def strip(m):
                if hasattr(m, 'weight_orig'):
                    remove_spectral_norm(m)
            self.averaged_model.apply(strip)
        else:
            self.averaged_model.eval()
        self._pairs = None

    def forward(self, *inputs, **kwargs):
        return self.module(*inputs, **kwargs)

    def _load_from_state_dict(self, *args, **kwargs):
        super()._load_from_state_dict(*args, **kwargs)
        self._steps = None      # resume: the warm-up counter continues from the checkpointed buffer

    @staticmethod
    def sn_compute_weight(weight, u, v):
        """W / sigma with the current u, v.  ref: model_average.py:183-197"""
        mat = weight.reshape(weight.shape[0], -1)
        sigma = torch.sum(u * torch.mv(mat, v))
        return weight / sigma

    def _build_pairs(self):
        src = self.module.state_dict()
        tgt = self.averaged_model.state_dict()
        plain_t, plain_s, sn = [], [], []
        for key, t in tgt.items():
            if key.endswith('weight') and key + '_orig' in src:
                sn.append((t, src[key + '_orig'], src[key + '_u'], src[key + '_v']))
            elif t.is_floating_point():
                plain_t.append(t)
                plain_s.append(src[key])
            else:
                sn.append((t, src[key], None, None))   # integer buffers: plain copy
        self._pairs = (plain_t, plain_s, sn)

    def _sync_steps(self):
        if self._steps is None:
            self._steps = int(self.num_updates_tracked.item())

    def next_beta(self):
        """beta the NEXT update_average() will use (utils/cuda_graph.py keys its captured iteration on it)."""
        self._sync_steps()
        return 0.0 if self._steps + 1 <= self.start_iteration else self.beta

    def note_graph_replay(self):
        """A captured update_average() has run on the device: advance the host mirror of num_updates_tracked."""
        self._sync_steps()
        self._steps += 1
        if not self.num_updates_tracked.is_cuda:
            # the reference builds ModelAverage after the network went to the GPU and never moves it (utils/trainer.py
            # :93-141): the counter is a host tensor, and a captured `+= 1` on it ran only at capture time
            self.num_updates_tracked += 1

    @torch.no_grad()
    def update_average(self):
        self._sync_steps()
        self.num_updates_tracked += 1
        self._steps += 1
        beta = 0.0 if self._steps <= self.start_iteration else self.beta
        if self._pairs is None:
            self._build_pairs()
        plain_t, plain_s, sn = self._pairs
        # target = beta * target + (1 - beta) * source
        torch._foreach_mul_(plain_t, beta)
        torch._foreach_add_(plain_t, plain_s, alpha=1.0 - beta)
        sn_layers = [e for e in sn if e[2] is not None]
        for t, w, u, v in sn:
            if u is None:
                t.copy_(w)
        if not sn_layers:
            return
        if not sn_layers[0][1].is_cuda:
            raise RuntimeError('ModelAverage.update_average runs on CUDA modules only (no CPU fallback)')
        # sigma = u^T W v of every spectral layer in one batched call (no power iteration), then
        # target = beta * target + (1 - beta) * W / sigma.   ref: model_average.py:183-197
        self._batched_sn_update(sn_layers, beta)

    def _batched_sn_update(self, sn_layers, beta):
        from .. import ops
        plan = self.__dict__.get('_ema_sn_plan')
        key = tuple((t.data_ptr(), w.data_ptr(), u.data_ptr(), v.data_ptr()) for t, w, u, v in sn_layers)
        if plan is None or plan['key'] != key:
            dev = sn_layers[0][1].device
            rows, t_off, out_off = [], 0, 0
            for _, w, u, v in sn_layers:
                r = w.shape[0]
                c = w.numel() // r
                rows.append([w.data_ptr(), u.data_ptr(), v.data_ptr(), t_off, 0, out_off, r, c])
                t_off += c
                out_off += 1 + r + c
            s_off = t_off
            for row in rows:
                row[4] = s_off
                s_off += row[6]
            plan = dict(key=key, descs=torch.tensor(rows, dtype=torch.int64).to(dev),
                        scratch=torch.empty((s_off,), dtype=torch.float32, device=dev), out_floats=out_off,
                        targets=torch.tensor([e[0].data_ptr() for e in sn_layers], dtype=torch.int64).to(dev),
                        max_elems=max(r[6] * r[7] for r in rows), max_r=max(r[6] for r in rows),
                        max_c=max(r[7] for r in rows))
            self.__dict__['_ema_sn_plan'] = plan
        out = ops.spectral_norm_sigma_batched(plan['descs'], len(sn_layers), plan['max_r'], plan['max_c'], plan['scratch'],
                                              plan['out_floats'], 0, 1e-12)
        ops.spectral_norm_ema_update(plan['descs'], plan['targets'], len(sn_layers), plan['max_elems'], out, beta)
