"""Attributes the ATen (non-library) kernels of one training iteration to Python source lines (developer tool)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import profile, ProfilerActivity
import bench
from imaginaire_b200.config import Config
from imaginaire_b200.trainers.spade import Trainer
from imaginaire_b200.utils.trainer import get_model_optimizer_and_scheduler

cfg = Config(os.path.join(bench.ROOT, 'configs', bench.CONFIG_NAME))
dev = torch.device('cuda:0')
net_G, net_D, opt_G, opt_D = get_model_optimizer_and_scheduler(cfg, seed=0, device=dev)
trainer = Trainer(cfg, net_G, net_D, opt_G, opt_D)
data = {k: v.to(dev) for k, v in bench.synthetic_batch(cfg, 8, 100).items()}
for _ in range(2):
    trainer.step(dict(data))
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA], with_stack=True) as prof:
    trainer.step(dict(data))
    torch.cuda.synchronize()
rows = []
for e in prof.key_averages(group_by_stack_n=6):
    t = getattr(e, 'self_device_time_total', 0)
    if e.key.startswith('aten::') and t > 50:
        stack = [s for s in e.stack if '/root/repo' in s or 'imaginaire_b200' in s][:3]
        rows.append((t, e.count, e.key, ' <- '.join(s.split('/')[-1] for s in stack)))
rows.sort(reverse=True)
for t, c, k, st in rows[:60]:
    print('%7.2f ms x%-4d %-28s %s' % (t / 1e3, c, k, st[:150]))
