"""Host-side logic of the N > 1 path with world_size 2 on the gloo backend (CPU only).
The data path has no collective besides the SyncBatchNorm statistic exchange and DDP's
gradient all-reduce; the statistic merge is plain torch code and is tested here."""
import os
import subprocess
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from tests import helpers as H


def _worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    from imaginaire_b200 import functional as NF
    from imaginaire_b200.utils import distributed as D
    D.init_dist(rank, backend='gloo')
    assert D.get_world_size() == world and D.get_rank() == rank and D.is_master() == (rank == 0)
    g = torch.Generator().manual_seed(123)
    full = torch.randn(5, 16, 6, 7, generator=g) * 1.3 + 0.4     # the "global batch"
    mine = full[:2] if rank == 0 else full[2:]                    # unequal per-rank batches
    sums = torch.stack([mine.sum((0, 2, 3)), (mine * mine).sum((0, 2, 3))]).view(2, 1, 16)
    count = mine.numel() // 16
    mean, rstd, unbiased = NF.sync_batch_stats(sums, count, 1e-5)
    ref_mean = full.mean((0, 2, 3))
    ref_var = full.var((0, 2, 3), unbiased=False)
    assert torch.allclose(mean, ref_mean, atol=1e-5)
    assert torch.allclose(rstd, (ref_var + 1e-5).rsqrt(), rtol=1e-4)
    assert torch.allclose(unbiased, full.var((0, 2, 3), unbiased=True), rtol=1e-4)
    # helper collectives (imaginaire/utils/distributed.py:61-93)
    t = torch.full((3,), float(rank + 1))
    assert torch.allclose(D.dist_all_reduce_tensor(t.clone()), torch.full((3,), 1.5))
    gathered = D.dist_all_gather_tensor(t.clone())
    assert [float(x[0]) for x in gathered] == [1.0, 2.0]
    # DDP over the reference-API blocks: parameters register and gradients average across ranks
    from torch.nn.parallel import DistributedDataParallel as DDP
    from imaginaire_b200.layers import LinearBlock
    torch.manual_seed(0)
    blk = DDP(LinearBlock(8, 4, weight_norm_type='spectral', nonlinearity='relu'))
    x = torch.full((2, 8), float(rank + 1))
    blk(x).sum().backward()
    grad = blk.module.layers['conv'].weight_orig.grad.clone()
    both = D.dist_all_gather_tensor(grad)
    assert torch.allclose(both[0], both[1])
    # the flat all-reduce that replaces a DDP wrapper for the small parameters (grad_sync.FlatGradSync) and the
    # initial state broadcast: gradients of the trainable parameters are averaged, frozen ones are left alone
    from imaginaire_b200 import grad_sync
    torch.manual_seed(rank)                         # different initial weights per rank on purpose
    net = torch.nn.Sequential(torch.nn.Linear(6, 5), torch.nn.BatchNorm1d(5), torch.nn.Linear(5, 3))
    grad_sync.broadcast_state(net)
    w0 = D.dist_all_gather_tensor(net[0].weight.detach().clone())
    assert torch.equal(w0[0], w0[1])
    flat = grad_sync.FlatGradSync(list(net.parameters()))
    net[2].bias.requires_grad_(False)
    net(torch.full((4, 6), float(rank + 1)) + torch.arange(4.0).view(4, 1)).square().sum().backward()
    local = [p.grad.clone() if p.grad is not None else None for p in net.parameters()]
    grad_sync.finish_all()
    for p, g in zip(net.parameters(), local):
        if g is None:
            assert p.grad is None
            continue
        both = D.dist_all_gather_tensor(g)
        assert torch.allclose(p.grad, (both[0] + both[1]) / 2, atol=1e-6)
    flat.close()
    with open(os.path.join(out_dir, 'ok%d' % rank), 'w') as f:
        f.write('ok')
    dist.destroy_process_group()


def test_world_size_2_gloo(tmp_path):
    port = 29500 + (os.getpid() % 2000)
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    assert (tmp_path / 'ok0').exists() and (tmp_path / 'ok1').exists()


def test_reference_arm_runs_on_rank0_only():
    """Under torchrun the CPU reference arm runs on rank 0; other ranks exit 0 without work."""
    env = dict(os.environ, RANK='1', WORLD_SIZE='2', LOCAL_RANK='1')
    r = subprocess.run([sys.executable, os.path.join(H.ROOT, 'bench.py'), '--impl', 'reference', '--gpus', '2',
                        '--steps', '1', '--warmup', '1'], env=env, capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and r.stdout.strip() == ''
