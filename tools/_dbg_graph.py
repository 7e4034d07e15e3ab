import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from imaginaire_b200.config import Config
from imaginaire_b200.trainers.spade import Trainer
from imaginaire_b200.utils.trainer import get_model_optimizer_and_scheduler
cfg = Config(os.path.join(bench.ROOT, 'configs', bench.CONFIG_NAME))
dev = torch.device('cuda', 0)
net_G, net_D, opt_G, opt_D, sch_G, sch_D = get_model_optimizer_and_scheduler(cfg, seed=0, device=dev)
trainer = Trainer(cfg, net_G, net_D, opt_G, opt_D, sch_G, sch_D, (), None)
net_G.train(); net_D.train()
static = {k: v.to(dev) for k, v in bench.synthetic_batch(cfg, 8, 100).items()}
def step():
    return trainer.step({k: v.detach() for k, v in static.items()})
def timeit(fn, n=10):
    torch.cuda.synchronize(); s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(n): fn()
    e.record(); torch.cuda.synchronize(); return s.elapsed_time(e) / n
for _ in range(4): step()
print('eager ms/step', timeit(step))
side = torch.cuda.Stream()
side.wait_stream(torch.cuda.current_stream())
with torch.cuda.stream(side):
    for _ in range(2): step()
torch.cuda.current_stream().wait_stream(side)
torch.cuda.synchronize()
g = torch.cuda.CUDAGraph()
try:
    with torch.cuda.graph(g):
        out = step()
    torch.cuda.synchronize()
    print('captured; losses', [float(o) for o in out])
    print('graph ms/step', timeit(g.replay))
    print('losses after replays', [float(o) for o in out])
except Exception as ex:
    import traceback; traceback.print_exc()
