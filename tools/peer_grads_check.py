"""Prints gradient checksums of the first training iteration (developer tool): run once with and once
without IMX_PEER_GRADS=1 under torchrun and compare - the peer-memory path must reproduce DDP's
averaged gradients up to fp32 summation order."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import bench
from imaginaire_b200.config import Config
from imaginaire_b200.trainers.spade import Trainer
from imaginaire_b200.utils.trainer import get_model_optimizer_and_scheduler

world, rank, local = int(os.environ.get('WORLD_SIZE', 1)), int(os.environ.get('RANK', 0)), int(os.environ.get('LOCAL_RANK', 0))
torch.cuda.set_device(local)
dev = torch.device('cuda', local)
if world > 1:
    dist.init_process_group('nccl', init_method='env://')
cfg = Config(os.path.join(bench.ROOT, 'configs', bench.CONFIG_NAME))
net_G, net_D, opt_G, opt_D, sch_G, sch_D = get_model_optimizer_and_scheduler(cfg, seed=0, device=dev)
trainer = Trainer(cfg, net_G, net_D, opt_G, opt_D, sch_G, sch_D, (), None)
data = {k: v.to(dev) for k, v in bench.synthetic_batch(cfg, 8, 100 + rank).items()}


def report(tag, net):
    def step():
        ps = [p for p in net.parameters() if p.grad is not None]
        big = [p for p in ps if p.dim() == 4 and p.numel() > 100000]
        small = [p for p in ps if not (p.dim() == 4 and p.numel() > 100000)]
        for name, group in (('filters', big), ('rest', small)):
            a = sum(float(p.grad.double().abs().sum()) for p in group)
            q = sum(float((p.grad.double() ** 2).sum()) for p in group)
            w = sum(float((p.grad.double() * torch.arange(p.numel(), device=p.device, dtype=torch.float64).view(p.shape).remainder(7.0)).sum()) for p in group)
            if rank == 0:
                print('%s %-8s n=%d  sum|g|=%.8e  sum g^2=%.8e  weighted=%.8e' % (tag, name, len(group), a, q, w), flush=True)
    return step


opt_D.step = report('D', net_D)
opt_G.step = report('G', net_G)
trainer.step(dict(data))
torch.cuda.synchronize()
