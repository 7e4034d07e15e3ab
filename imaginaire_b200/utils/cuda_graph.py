"""One training iteration as a CUDA graph.

A SPADE G+D iteration is ~3 900 kernel launches, many of them 5-20 us long; issued one by one from Python they leave
the GPU idle between launches (measured: 134.4 ms per iteration eager, 128.9 ms replayed, 127 ms of kernel time).
``GraphedTrainerStep`` captures ``trainer.dis_update`` + ``trainer.gen_update`` (train.py:76-79's order) once and
replays them.  What makes the replay equal to the eager loop:

  * inputs live in static device tensors that each call copies the new batch into (``copy_`` advances their version
    counters, so per-batch caches kept on the label tensor were rebuilt at capture time and the rebuild is part of
    the graph);
  * Adam's step numbers are device counters (``FusedAdam.make_capturable``), the kernel forms the bias corrections;
  * values that change rarely and are baked into kernel arguments -- the EMA's beta (0 during the warm-up of
    ``model_average_start_iteration`` updates, then ``model_average_beta``) and the learning rates -- form the capture
    key: when the next iteration's key differs, the iteration is captured again;
  * host-side counters that the captured Python body would have advanced (the EMA's warm-up count) are advanced here.

Random draws (the style encoder's noise) go through torch's graph-safe CUDA generator.  Data-parallel runs: every
rank captures and replays in step; what crosses ranks inside the iteration is kernels only (the peer-memory
exchanges of peer_exchange.py / grad_sync.py with device-side counters, symmetric-memory barriers, one NCCL
all-reduce of the flat small-parameter gradients), all of which a capture records.  A DistributedDataParallel
wrapper (host-side hooks and buckets) keeps the eager loop.  Run at least one eager iteration first (lazy set-up: kernel
attributes, spectral-norm plans, optimizer state); ``warmup=n`` runs n more eager iterations on the first batch.
"""
import torch


class GraphedTrainerStep:
    def __init__(self, trainer, example_data, warmup=0):
        self.trainer = trainer
        self.static = {k: torch.empty_like(v) for k, v in example_data.items() if isinstance(v, torch.Tensor)}
        self.graph = None
        self.key = None
        self.warmup = warmup
        self.launches_per_replay = 0
        self._params = [p for opt in (trainer.opt_G, trainer.opt_D) for g in opt.param_groups for p in g['params']]
        for opt in (trainer.opt_G, trainer.opt_D):
            if hasattr(opt, 'make_capturable'):
                opt.make_capturable()
            else:
                raise RuntimeError('GraphedTrainerStep needs the library FusedAdam (device-side step counters)')

    # values baked into the captured kernel arguments
    def _key(self):
        t = self.trainer
        ema = t.net_G.module if t.cfg.trainer.model_average else None
        beta = ema.next_beta() if ema is not None else None
        return (beta, tuple(g['lr'] for g in t.opt_G.param_groups), tuple(g['lr'] for g in t.opt_D.param_groups))

    def _body(self):
        # shallow copies: the generator adds entries (its style code) to the dictionary it is given
        self.trainer.dis_update(dict(self.static))
        self.trainer.gen_update(dict(self.static))

    def _capture(self):
        from .. import _C
        ema = self.trainer.net_G.module if self.trainer.cfg.trainer.model_average else None
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(self.warmup if self.graph is None else 0):
                self._body()              # eager iterations: lazy set-up (kernel attributes, plans, caches) happens here
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        # Packed bf16 filters are cached per parameter version.  An entry left by the eager iterations would be a cache
        # HIT during capture: no pack kernel would be recorded, every replay would read that one tensor, never
        # refreshed -- and freed for reuse as soon as the cache drops it.  Start the capture with an empty cache.
        from .. import functional as NF
        NF._PACK_CACHE.clear()
        lib = _C.lib()
        before = lib.launches
        steps_before = ema._steps if ema is not None else None
        tracked_before = ema.num_updates_tracked.clone() if ema is not None and not ema.num_updates_tracked.is_cuda else None
        self.graph = torch.cuda.CUDAGraph()
        # thread_local: the NCCL watchdog thread of a data-parallel run keeps querying its events while this thread
        # captures; under the default (global) mode those queries are capture violations
        with torch.cuda.graph(self.graph, capture_error_mode='thread_local'):
            self._body()
        self.launches_per_replay = lib.launches - before
        if ema is not None:
            ema._steps = steps_before     # capture runs the Python body once without executing a kernel
            if tracked_before is not None:
                ema.num_updates_tracked.copy_(tracked_before)
        self.key = self._key()

    def close(self):
        """Drops the captured graph.  Call before torch.distributed.destroy_process_group(): a graph that recorded NCCL
        work must not outlive its communicator."""
        torch.cuda.synchronize()
        self.graph = None
        self.key = None

    def __call__(self, data):
        """Runs one iteration on `data` (device or pinned-host tensors); returns (generator total, discriminator
        total) like Trainer.step.  The returned tensors are the graph's static outputs: read them before the next call."""
        key = self._key()
        if self.graph is None:
            for k, v in self.static.items():
                v.copy_(data[k], non_blocking=True)
            self._capture()
            key = self.key
        elif key != self.key:
            self._capture()
        for k, v in self.static.items():
            v.copy_(data[k], non_blocking=True)
        self.graph.replay()
        # the captured optimizer steps rewrite the parameters on the device; tell the host-side caches that key on
        # Tensor._version (packed bf16 filters, identity-affine checks) that their values moved
        torch._C._increment_version(self._params)
        t = self.trainer
        if t.cfg.trainer.model_average:
            t.net_G.module.note_graph_replay()
        from .. import _C
        _C.lib().launches += self.launches_per_replay
        return t.gen_losses['total'], t.dis_losses['total']
