"""Data-parallel filter gradients over NVLink peer memory, fused with the split-K reduction.

ref: the reference wraps its networks in (apex / torch) DistributedDataParallel, whose bucketed
NCCL all-reduce averages the gradients after the backward (imaginaire/utils/trainer.py:193-216).
Here the convolution filters (99 % of the parameters) take another route: the weight-gradient
GEMM already ends in a reduction over split-K partial sums; with every rank's partials in
symmetric (peer-mapped) memory that reduction simply also runs over the ranks (``IMX_PEER_GRADS_MODE=partials``):

    per layer:    GEMM -> partials (own region of a symmetric workspace) imx_conv2d_igemm_wgrad_partials_bf16
    per backward: barrier (device side, current stream)
                 per layer: sum over ranks x splits of MY 1/world of the channels  imx_wgrad_reduce_peers_f32
                 barrier, fetch the other ranks' channel ranges        imx_gather_rows_peers_f32
                 spectral-norm gradient correction on the averaged gradients (it is linear)

That moves `splits` x the gradient bytes over NVLink (7.5 ms of reduction kernels per iteration on 2 GPUs, ncu /
torch.profiler).  The default (``IMX_PEER_GRADS_MODE=flat``) keeps the split-K reduction local and exchanges finished
gradients only:

    per layer:    the ordinary weight gradient (GEMM + local split-K reduction) written into the rank's arena
    per backward: barrier; ONE reduce-scatter kernel over all gradients of the backward (imx_peer_reduce_scatter_f32:
                 my 1/world chunk of every tensor summed over the ranks); barrier; ONE all-gather kernel
                 (imx_peer_all_gather_f32); barrier; spectral-norm correction

No NCCL kernels compete with the persistent convolution kernels for SMs, and the averaged
gradient is produced by kernels that had to run anyway.  Everything else (biases, norm
parameters, linear layers, layers with <= 8 channels) stays with DDP.

On by default for world sizes 2..8 on one node (utils/trainer.py); ``IMX_PEER_GRADS=0`` leaves every
parameter to DDP.  Measured (bench.py, C3 step): 2 GPUs 159 -> 155 ms per step; 8 GPUs 168 -> 169 ms
(no gain: there the step is bound by the 69 latency-serialised SyncBN exchanges and the slowest GPU).
"""
import ctypes

import torch
import torch.distributed as dist

from . import _C, ops

_ACTIVE = {}      # data_ptr of a registered filter -> PeerGradSync


def lookup(weight):
    return _ACTIVE.get(weight.data_ptr())


class PeerGradSync:
    def __init__(self, weights, workspace_bytes=6 << 30, group=None, max_uses=1, extra_params=None):
        import torch.distributed._symmetric_memory as symm_mem
        self.group = group if group is not None else dist.group.WORLD
        self.world = dist.get_world_size(self.group)
        self.rank = dist.get_rank(self.group)
        if not 2 <= self.world <= 8:
            raise RuntimeError('PeerGradSync supports 2..8 ranks on one node (got %d)' % self.world)
        dev = weights[0].device
        self.params = list(weights)
        # One gradient slot per (filter, use): a network that is applied several times in one forward
        # (the discriminator on real and on fake images) runs a spectral-norm power iteration per
        # application, so each application's gradient needs its own (u, v, sigma) correction before
        # the applications are summed.
        self.max_uses = max_uses
        self.offset = {}
        total = 0
        for use in range(max_uses):
            for w in self.params:
                self.offset[(w.data_ptr(), use)] = total
                total += (w.numel() + 63) // 64 * 64          # 256-byte aligned slots
        # `extra_params`: everything else that needs averaging (biases, norm parameters, linear layers, filters with
        # <= 8 channels).  Their gradients are gathered into one more segment at the end of the arena and ride the same
        # reduce-scatter / all-gather launches and barriers: no NCCL call is left in the iteration.
        self.extra_params = list(extra_params) if extra_params is not None else []
        self.extra_off = total
        total += (sum(p.numel() for p in self.extra_params) + 63) // 64 * 64
        self.arena = symm_mem.empty(total, dtype=torch.float32, device=dev)
        self.arena_hdl = symm_mem.rendezvous(self.arena, self.group.group_name)
        import os
        self.flat = os.environ.get('IMX_PEER_GRADS_MODE', 'flat') != 'partials'
        self.arena_ptrs = [int(p) for p in self.arena_hdl.buffer_ptrs]
        if not self.flat:      # the split-K partial sums are exchanged: they need a symmetric workspace
            self.ws = symm_mem.empty(workspace_bytes // 4, dtype=torch.float32, device=dev)
            self.ws_hdl = symm_mem.rendezvous(self.ws, self.group.group_name)
            self.ws_ptrs = [int(p) for p in self.ws_hdl.buffer_ptrs]
        self.ws_bytes = workspace_bytes
        self.ws_regions = {}      # (data_ptr, use index) -> byte offset; same order on every rank
        self.ws_top = 0
        self.uses = {}            # data_ptr -> times used in this backward
        self.pending_sn = []
        self.records = []
        self.by_ptr = {w.data_ptr(): w for w in self.params}
        self._gather_descs = {}
        self._seg_descs = {}
        self._arena_table = (ctypes.c_void_p * self.world)(*self.arena_ptrs)
        for w in self.params:
            _ACTIVE[w.data_ptr()] = self

    def close(self):
        """Unregisters the filters (their gradients go back to whatever data-parallel wrapper is used)."""
        for w in self.params:
            _ACTIVE.pop(w.data_ptr(), None)

    # ------------------------------------------------------------------ per layer
    def _region(self, key, nbytes):
        off = self.ws_regions.get(key)
        if off is None:
            off = self.ws_top
            self.ws_top += (nbytes + 255) // 256 * 256
            if self.ws_top > self.ws_bytes:
                raise RuntimeError('PeerGradSync: split-K workspace arena exhausted (%d bytes); raise IMX_PEER_GRADS_WS_GB'
                                   % self.ws_bytes)
            self.ws_regions[key] = off
        return off

    def grad_view(self, weight, use=0):
        off = self.offset[(weight.data_ptr(), use)]
        return self.arena[off:off + weight.numel()].view(weight.shape)

    def wgrad(self, weight, x, dy, cin, cout, kh, kw, stride, pad, sn=None):
        """Runs the weight-gradient GEMM and this rank's share of the cross-rank reduction.  The
        complete averaged gradient is in grad_view(weight) after finish()."""
        lib = _C.lib()
        n, cin_p, h, w = x.shape
        cout_p = dy.shape[1]
        nbytes = lib.raw('imx_conv2d_wgrad_workspace_bytes')(n, h, w, cin, cin_p, cout, cout_p, kh, kw, stride, pad)
        if nbytes < 0:
            raise _C.ImxError('imx_conv2d_wgrad_workspace_bytes: %s' % lib.last_error())
        key = weight.data_ptr()
        use = self.uses.get(key, 0)
        if use >= self.max_uses:
            raise RuntimeError('PeerGradSync: a filter is used %d times in one backward but max_uses is %d'
                               % (use + 1, self.max_uses))
        if self.flat:
            # the complete local gradient (GEMM + local split-K reduction + this application's spectral-norm
            # correction, which is linear in the gradient: correcting before the average equals correcting after, and
            # nothing order-dependent is left to do once the ranks hold the same bits) straight into this rank's slot
            ops.conv2d_wgrad(x, dy, cin, cout, kh, kw, stride, pad, out=self.grad_view(weight, use), sn=sn)
            self.uses[key] = use + 1
            return
        off = self._region((key, use), int(nbytes))
        plan = (ctypes.c_int * 3)()
        stream = ops._stream()
        lib.call('imx_conv2d_igemm_wgrad_partials_bf16', x.data_ptr(), dy.data_ptr(), self.ws_ptrs[self.rank] + off,
                 int(nbytes), plan, n, h, w, cin, cin_p, cout, cout_p, kh, kw, stride, pad, stream)
        # the cross-rank reduction of ALL layers runs in finish(), behind one barrier: a barrier per
        # layer would add ~160 cross-GPU synchronisation points to every backward
        self.records.append((key, use, off, int(plan[0]), int(plan[1]), int(plan[2]), cin, cout, kh * kw))
        self.uses[key] = use + 1
        if sn is not None:
            self.pending_sn.append((key, use) + tuple(sn))

    # ------------------------------------------------------------------ per backward
    def finish(self):
        """Call after backward() and before optimizer.step(): completes the gradients of every filter
        that took part in this backward and installs them as ``.grad``."""
        # a registered filter whose gradient came through plain autograd (used as W / sigma by torch
        # code instead of the native convolution) is averaged the ordinary way
        for k, p in self.by_ptr.items():
            if k not in self.uses and p.grad is not None and p.grad.data_ptr() != self.grad_view(p).data_ptr():
                dist.all_reduce(p.grad, op=dist.ReduceOp.SUM, group=self.group)
                p.grad.div_(self.world)
        extra = [p.grad for p in self.extra_params if p.requires_grad and p.grad is not None] if self.flat else []
        if not self.uses and not extra:
            return
        used = tuple(sorted(((k, u) for k, n in self.uses.items() for u in range(n)), key=lambda ku: self.offset[ku]))
        descs = self._gather_descs.get(used)
        if descs is None:
            rows = [[self.offset[ku], self.by_ptr[ku[0]].shape[0], self.by_ptr[ku[0]].numel() // self.by_ptr[ku[0]].shape[0]]
                    for ku in used]
            descs = (torch.tensor(rows, dtype=torch.int64).to(self.arena.device), max(self.by_ptr[k].numel() for k, _ in used))
            self._gather_descs[used] = descs
        lib = _C.lib()
        stream = ops._stream()
        if self.flat:
            n_extra = sum(g.numel() for g in extra)
            extra_views = None
            if n_extra:
                flat = self.arena[self.extra_off:self.extra_off + n_extra]
                torch.cat([g.reshape(-1) for g in extra], out=flat)
                extra_views = [c.view_as(g) for c, g in zip(flat.split([g.numel() for g in extra]), extra)]
            segs = self._seg_descs.get((used, n_extra))
            if segs is None:
                rows = [[self.offset[ku], self.by_ptr[ku[0]].numel()] for ku in used]
                if n_extra:
                    rows.append([self.extra_off, n_extra])
                segs = (torch.tensor(rows, dtype=torch.int64).to(self.arena.device), max(r[1] for r in rows), len(rows))
                self._seg_descs[(used, n_extra)] = segs
            n_seg = segs[2]
            self.arena_hdl.barrier(channel=0)              # every rank's local gradients of this backward are written
            lib.call('imx_peer_reduce_scatter_f32', self._arena_table, self.world, self.rank, segs[0].data_ptr(), n_seg,
                     int(segs[1]), 1.0 / self.world, stream)
            self.arena_hdl.barrier(channel=1)              # every rank has reduced its chunks
            lib.call('imx_peer_all_gather_f32', self._arena_table, self.world, self.rank, segs[0].data_ptr(), n_seg,
                     int(segs[1]), stream)
            self.arena_hdl.barrier(channel=2)              # nobody reads my chunks any more: in-place edits are safe
            if extra_views is not None:
                torch._foreach_copy_(extra, extra_views)
            self._finish_local()
            return
        self.ws_hdl.barrier(channel=0)                     # every rank's partial sums of this backward are written
        for key, use, off, splits, co_r, ci_r, cin, cout, taps in self.records:
            peers = (ctypes.c_void_p * self.world)(*[p + off for p in self.ws_ptrs])
            dw = self.grad_view(self.by_ptr[key], use)
            co_b, co_e = self.rank * cout // self.world, (self.rank + 1) * cout // self.world
            lib.call('imx_wgrad_reduce_peers_f32', peers, self.world, dw.data_ptr(), co_b, co_e, cout, cin, taps,
                     splits, co_r, ci_r, 1.0 / self.world, 0, stream)
        self.records.clear()
        self.arena_hdl.barrier(channel=1)                  # every rank has reduced its channel ranges
        peers = (ctypes.c_void_p * self.world)(*self.arena_ptrs)
        lib.call('imx_gather_rows_peers_f32', peers, self.world, self.rank, descs[0].data_ptr(), len(used), int(descs[1]),
                 ops._stream())
        self.arena_hdl.barrier(channel=2)                  # nobody reads my channel ranges any more: in-place edits are safe
        self._finish_local()

    def _finish_local(self):
        for key, use, weight, u, v, inv_sigma in self.pending_sn:
            ops.spectral_norm_weight_grad_(self.grad_view(self.by_ptr[key], use), weight, u, v, inv_sigma)
        for k, n in self.uses.items():
            p = self.by_ptr[k]
            g = self.grad_view(p)
            for use in range(1, n):
                g.add_(self.grad_view(p, use))
            p.grad = g
        self.uses.clear()
        self.pending_sn.clear()


class FlatGradSync:
    """Averages the gradients of everything PeerGradSync does not carry (biases, norm parameters, linear layers,
    filters with <= 8 channels on one side: ~1 % of the parameters) with ONE all-reduce of a flat buffer after the
    backward: gather into the buffer (torch.cat), NCCL all-reduce, scatter back (_foreach_copy_).  It replaces the
    DistributedDataParallel wrapper for these parameters: no autograd hooks, no bucket bookkeeping, no
    unused-parameter search, and nothing that a CUDA-graph capture of the iteration cannot record.
    ref: the gradient averaging of the reference's DDP wrapper, imaginaire/utils/trainer.py:193-216."""

    def __init__(self, params, group=None):
        self.group = group if group is not None else dist.group.WORLD
        self.world = dist.get_world_size(self.group)
        self.params = [p for p in params]
        _FLAT.append(self)

    def close(self):
        if self in _FLAT:
            _FLAT.remove(self)

    def finish(self):
        grads = [p.grad for p in self.params if p.grad is not None]
        if not grads:
            return
        flat = torch.cat([g.reshape(-1) for g in grads])
        dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=self.group)
        flat.mul_(1.0 / self.world)
        torch._foreach_copy_(grads, [c.view_as(g) for c, g in zip(flat.split([g.numel() for g in grads]), grads)])


_FLAT = []


def broadcast_state(model, group=None, src=0):
    """Rank `src`'s parameters and buffers to every rank (what the DDP constructor does), coalesced per dtype and
    device.  Host-side tensors (the EMA wrapper's update counter) are skipped under NCCL, which moves device memory
    only: they are counters that every rank advances identically."""
    by_kind = {}
    nccl = dist.get_backend(group) == 'nccl'
    for t in list(model.parameters()) + list(model.buffers()):
        if nccl and not t.is_cuda:
            continue
        by_kind.setdefault((t.dtype, t.device), []).append(t)
    with torch.no_grad():
        for ts in by_kind.values():
            flat = torch.cat([t.reshape(-1) for t in ts])
            dist.broadcast(flat, src=src, group=group)
            torch._foreach_copy_(ts, [c.view_as(t) for c, t in zip(flat.split([t.numel() for t in ts]), ts)])


def finish_all():
    for s in {id(v): v for v in _ACTIVE.values()}.values():
        s.finish()
    for f in _FLAT:
        f.finish()


def eligible_filters(model):
    """(name, parameter) of the convolution filters whose gradient GEMM goes through the generic
    tcgen05 weight-gradient kernel: more than 8 channels on both sides (the small-channel layers
    take the im2col / tap-sum routes of functional.py and stay with DDP)."""
    from .layers.conv import Conv2d
    out = []
    for mod_name, m in model.named_modules():
        if isinstance(m, Conv2d) and m.in_channels > 8 and m.out_channels > 8:
            pname = 'weight_orig' if hasattr(m, 'weight_orig') else 'weight'
            p = getattr(m, pname)
            if isinstance(p, torch.nn.Parameter) and p.requires_grad and p.is_cuda:
                out.append(((mod_name + '.' if mod_name else '') + pname, p))
    return out
