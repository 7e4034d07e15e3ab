"""Host-side enqueue time of one training iteration vs its GPU time (developer tool)."""
import os, sys, time, cProfile, pstats, io
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from imaginaire_b200.config import Config
from imaginaire_b200.trainers.spade import Trainer
from imaginaire_b200.utils.trainer import get_model_optimizer_and_scheduler

cfg = Config(os.path.join(bench.ROOT, 'configs', bench.CONFIG_NAME))
dev = torch.device('cuda:0')
net_G, net_D, opt_G, opt_D, sch_G, sch_D = get_model_optimizer_and_scheduler(cfg, seed=0, device=dev)
trainer = Trainer(cfg, net_G, net_D, opt_G, opt_D, sch_G, sch_D, (), None)
data = {k: v.to(dev) for k, v in bench.synthetic_batch(cfg, 8, 100).items()}
for _ in range(3):
    trainer.step(dict(data))
torch.cuda.synchronize()
for _ in range(3):
    t0 = time.perf_counter()
    trainer.step(dict(data))
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    print('enqueue %.1f ms, until GPU done %.1f ms' % ((t1 - t0) * 1e3, (t2 - t0) * 1e3))
pr = cProfile.Profile()
pr.enable()
trainer.step(dict(data))
pr.disable()
torch.cuda.synchronize()
s = io.StringIO()
pstats.Stats(pr, stream=s).sort_stats('tottime').print_stats(35)
print(s.getvalue()[:6000])
