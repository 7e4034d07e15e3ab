"""SyncBatchNorm's cross-rank exchanges over NVLink peer memory (csrc/peer_exchange.cu).

ref: torch.nn.SyncBatchNorm (torch/nn/modules/_functions.py:49-165), which the reference instantiates for
``sync_batch`` (imaginaire/layers/activation_norm.py:403-410): one NCCL collective per layer and pass.  A C3 iteration
runs 69 of them, each a few KB: pure latency on the compute stream.  ``PeerExchange`` keeps one region of symmetric
memory per rank and does each exchange in a single one-CTA kernel that also finalizes the statistics.

One exchange object per process group, created on first use (``get``); if symmetric memory cannot be set up on every
rank (not one node, no peer access) ``get`` returns None and the callers keep their NCCL path.
``IMX_PEER_SYNCBN=0`` forces that.
"""
import ctypes
import os

import torch
import torch.distributed as dist

from . import _C, ops

_EXCHANGES = {}


class PeerExchange:
    MAX_FLOATS = 8192

    def __init__(self, group=None, regions=None, rank=None):
        """regions / rank: for tests that emulate several ranks on one GPU with plain tensors."""
        lib = _C.lib()
        self.nbytes = int(lib.raw('imx_peer_exchange_region_bytes')())
        if regions is not None:
            self.world, self.rank = len(regions), rank
            self._keep = regions
            self.ptrs = [r.data_ptr() for r in regions]
        else:
            import torch.distributed._symmetric_memory as symm_mem
            self.group = group if group is not None else dist.group.WORLD
            self.world = dist.get_world_size(self.group)
            self.rank = dist.get_rank(self.group)
            if not 2 <= self.world <= 8:
                raise RuntimeError('PeerExchange supports 2..8 ranks on one node (got %d)' % self.world)
            self.region = symm_mem.empty(self.nbytes // 4, dtype=torch.float32, device=torch.device('cuda', torch.cuda.current_device()))
            self.region.zero_()
            self.hdl = symm_mem.rendezvous(self.region, self.group.group_name)
            self.ptrs = [int(p) for p in self.hdl.buffer_ptrs]
            torch.cuda.synchronize()
            self.hdl.barrier(channel=0)      # every region is zeroed before anybody's first push arrives
            torch.cuda.synchronize()
        self._table = (ctypes.c_void_p * self.world)(*self.ptrs)

    def allreduce_sum(self, vec):
        """Sum of `vec` (fp32, <= 8192 elements) over the ranks; a new tensor, identical on every rank."""
        flat = vec.reshape(-1)
        if not flat.is_contiguous():
            flat = flat.contiguous()
        out = torch.empty_like(flat)
        _C.lib().call('imx_peer_allreduce_sum_f32', self._table, self.world, self.rank, flat.data_ptr(), out.data_ptr(),
                      flat.numel(), ops._stream())
        return out.view(vec.shape)

    def allreduce_sum_(self, vec):
        """In place."""
        if not vec.is_contiguous():
            raise RuntimeError('allreduce_sum_: contiguous tensor expected')
        _C.lib().call('imx_peer_allreduce_sum_f32', self._table, self.world, self.rank, vec.data_ptr(), vec.data_ptr(),
                      vec.numel(), ops._stream())
        return vec

    def bn_finalize(self, sums, count, eps, running_mean=None, running_var=None, momentum=0.1):
        """sums: (2, 1, C) of this rank; count: GLOBAL elements per channel.  Returns (mean, rstd), each (C,)."""
        c = sums.shape[-1]
        if not sums.is_contiguous():
            sums = sums.contiguous()
        out = torch.empty((2, c), dtype=torch.float32, device=sums.device)
        _C.lib().call('imx_peer_bn_finalize_f32', self._table, self.world, self.rank, sums.data_ptr(), c, float(count),
                      float(eps), out[0].data_ptr(), out[1].data_ptr(), ops._ptr(running_mean), ops._ptr(running_var),
                      float(momentum), ops._stream())
        return out[0], out[1]


def get(group=None):
    """The exchange of a process group, or None when the peer path is unavailable (all ranks agree)."""
    key = group if group is not None else 'world'
    if key in _EXCHANGES:
        return _EXCHANGES[key]
    ex = None
    if os.environ.get('IMX_PEER_SYNCBN', '1') != '0' and dist.is_available() and dist.is_initialized() \
            and dist.get_backend(group) == 'nccl' and 2 <= dist.get_world_size(group) <= 8:
        ok = torch.ones(1, device='cuda')
        try:
            ex = PeerExchange(group)
        except Exception as e:     # noqa: BLE001 - any failure means "keep NCCL"
            print('peer-memory SyncBN exchange unavailable (%s); using NCCL' % repr(e)[:200])
            ok.zero_()
        dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)
        if float(ok) == 0:
            ex = None
    _EXCHANGES[key] = ex
    return ex
