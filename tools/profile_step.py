"""Kernel-time breakdown of one training iteration with torch.profiler (developer tool)."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import profile, ProfilerActivity
import bench
from imaginaire_b200.config import Config
from imaginaire_b200.trainers.spade import Trainer
from imaginaire_b200.utils.trainer import get_model_optimizer_and_scheduler

cfg = Config(os.path.join(bench.ROOT, 'configs', bench.CONFIG_NAME))
world, rank, local = int(os.environ.get('WORLD_SIZE', 1)), int(os.environ.get('RANK', 0)), int(os.environ.get('LOCAL_RANK', 0))
torch.cuda.set_device(local)
dev = torch.device('cuda', local)
if world > 1:
    import torch.distributed as dist
    dist.init_process_group('nccl', init_method='env://')
net_G, net_D, opt_G, opt_D, sch_G, sch_D = get_model_optimizer_and_scheduler(cfg, seed=0, device=dev)
trainer = Trainer(cfg, net_G, net_D, opt_G, opt_D, sch_G, sch_D, (), None)
data = {k: v.to(dev) for k, v in bench.synthetic_batch(cfg, int(sys.argv[1]) if len(sys.argv) > 1 else 8, 100).items()}
for _ in range(2):
    trainer.step(dict(data))
torch.cuda.synchronize()
step = trainer.step
if os.environ.get('GRAPH'):        # GRAPH=1: profile the CUDA-graph replay (what bench.py times) instead of the eager loop
    from imaginaire_b200.utils.cuda_graph import GraphedTrainerStep
    step = GraphedTrainerStep(trainer, data)
    for _ in range(3):
        step(dict(data))
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
ATEN = bool(os.environ.get('ATEN'))     # ATEN=1: also list the ATen ops by their own kernel time, with shapes and call sites
_xc = torch._C._profiler._ExperimentalConfig(verbose=True) if ATEN else None     # needed for Python stacks in key_averages
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA], record_shapes=ATEN, with_stack=ATEN,
             experimental_config=_xc) as prof:
    step(dict(data))
    torch.cuda.synchronize()
if os.environ.get('GRAPH'):
    step.close()
if rank != 0:
    sys.exit(0)
if ATEN:
    ops_rows = []
    for e in prof.key_averages(group_by_input_shape=True, group_by_stack_n=12):
        t = getattr(e, 'self_device_time_total', 0)
        if e.key.startswith('aten::') and t > 0:
            where = [f for f in e.stack if 'imaginaire_b200' in f or 'bench.py' in f]
            ops_rows.append((t, e.count, e.key, str(e.input_shapes)[:90], (where[0] if where else (e.stack[0] if e.stack else ''))[-90:]))
    ops_rows.sort(reverse=True)
    print('ATen ops by own kernel time: %.1f ms total' % (sum(r[0] for r in ops_rows) / 1e3))
    for t, c, k, shp, where in ops_rows[:int(os.environ.get('TOP', 45))]:
        print('%7.3f ms x%-4d %-28s %-90s %s' % (t / 1e3, c, k, shp, where))
rows = []
for e in prof.key_averages():
    t = getattr(e, 'device_time_total', None)
    if t is None:
        t = getattr(e, 'cuda_time_total', 0)
    if e.device_type == torch.autograd.DeviceType.CUDA and t > 0:
        rows.append((t, e.count, e.key))
rows.sort(reverse=True)
tot = sum(r[0] for r in rows)
print('total device kernel time %.1f ms over %d kernel names' % (tot / 1e3, len(rows)))
for t, c, k in rows[:int(os.environ.get("TOP", 45))]:
    print('%8.2f ms %5.1f%% x%-5d %s' % (t / 1e3, 100.0 * t / tot, c, k[:110]))
