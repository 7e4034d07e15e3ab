"""Hardware check of the data-parallel path (run under torchrun on >= 2 GPUs of one node; developer tool).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
        tools/multi_gpu_check.py > profiles/r02_multi_gpu_check_n2.txt

1. SyncBatchNorm across ranks == BatchNorm of the concatenated batch on one device (SURVEY 7.2 step 8's gate; ref
   imaginaire/layers/activation_norm.py:403-410 + torch nn/modules/_functions.py:49-165): a SPADE norm + LeakyReLU with
   a sync_batch inner norm runs forward and backward on every rank's own quarter / half of a batch through the
   peer-memory exchange (csrc/peer_exchange.cu); rank 0 gathers the pieces and compares them with the fp32 CPU oracle on
   the whole batch (outputs, input gradients, running statistics).
2. Gradients of one training iteration: the peer-memory filter path + the flat all-reduce of the rest (grad_sync.py)
   against torch DistributedDataParallel over NCCL for everything, same networks, same per-rank batches
   (configs/spade_c3_mini.yaml), and both against ONE GPU on the concatenated batch (run first with `--single W`,
   which leaves its gradients in /tmp).  Printed per parameter group: relative L2 difference of the averaged gradients.
   The discriminator's gradients are well conditioned and agree to bf16 accuracy; the generator's early layers are
   not (profiles/r02_generator_grad_trace.md: bf16 storage of the activation gradients alone moves them by 25-35 %),
   and two runs of ANY variant differ there by as much as two variants do -- what the table has to show is that the
   three pairs differ by the same amount.
"""
import os
import sys
import types

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from imaginaire_b200 import functional as NF, grad_sync, ops, peer_exchange
from imaginaire_b200.config import Config
from oracle import spade_oracle as so
from tests import helpers as H

SINGLE = int(sys.argv[sys.argv.index('--single') + 1]) if '--single' in sys.argv else 0
SINGLE_FILE = '/tmp/imx_single_gpu_grads.pt'
world, rank, local = int(os.environ.get('WORLD_SIZE', 1)), int(os.environ.get('RANK', 0)), int(os.environ.get('LOCAL_RANK', 0))
torch.cuda.set_device(local)
dev = torch.device('cuda', local)
if not SINGLE:
    dist.init_process_group('nccl', init_method='env://')


def say(*a):
    if rank == 0:
        print(*a, flush=True)


def rel_l2(a, b):
    return float((a.double() - b.double()).norm() / (b.double().norm() + 1e-30))


# ------------------------------------------------------------------ 1. SyncBN == BN of the concatenated batch
def check_syncbn():
    from imaginaire_b200.layers import SpatiallyAdaptiveNorm
    torch.manual_seed(5)
    c, cond, n_local, h, w = 64, 6, 2, 24, 40
    m = SpatiallyAdaptiveNorm(c, cond, num_filters=16, kernel_size=3, separate_projection=True,
                              activation_norm_type='sync_batch', activation_norm_params=types.SimpleNamespace(affine=True))
    with torch.no_grad():
        m.norm.weight.uniform_(0.5, 1.5)
        m.norm.bias.uniform_(-0.5, 0.5)
    sd0 = {k: v.clone() for k, v in m.state_dict().items()}
    g = torch.Generator().manual_seed(6)
    x_all = torch.randn(world * n_local, c, h, w, generator=g) * 1.5 + 0.3
    x_all = x_all.bfloat16().float()                 # the kernels store x in bf16: compare on identical inputs
    lab_all = H.onehot_label(7, world * n_local, cond, 4 * h, 4 * w, blocky=4)
    go_all = torch.randn(world * n_local, c, h, w, generator=g)
    sl = slice(rank * n_local, (rank + 1) * n_local)
    m = m.to(dev).train()
    x = x_all[sl].to(dev).requires_grad_(True)
    y = NF.to_nchw(m.forward_fused(x, lab_all[sl].to(dev), act=ops.ACT_LRELU), c)
    y.backward(go_all[sl].to(dev))
    used = peer_exchange.get(None) is not None
    ys, gxs = [torch.empty_like(y) for _ in range(world)], [torch.empty_like(x) for _ in range(world)]
    dist.all_gather(ys, y.detach().contiguous())
    dist.all_gather(gxs, x.grad.contiguous())
    rm, rv = m.norm.running_mean.clone(), m.norm.running_var.clone()
    rms = [torch.empty_like(rm) for _ in range(world)]
    dist.all_gather(rms, rm)
    if rank == 0:
        xo = x_all.clone().requires_grad_(True)
        p = dict(kernel_size=3, separate_projection=True, num_filters=16, norm_kind='sync_batch')
        sd = {k: v.clone() for k, v in sd0.items()}
        yo = so.leaky_relu(so.spatially_adaptive_norm(sd, '', xo, [lab_all], p, True))
        yo.backward(go_all)
        print('1. SyncBN over %d ranks (peer-memory exchange: %s) vs BatchNorm of the concatenated batch (fp32 oracle)' % (world, used))
        print('   output  rel-L2 %.3e   (bf16 output storage: ~2e-3)' % rel_l2(torch.cat(ys).cpu(), yo.detach()))
        print('   dL/dx   rel-L2 %.3e' % rel_l2(torch.cat(gxs).cpu(), xo.grad))
        print('   running_mean rel %.3e  running_var rel %.3e' % (H.rel_err(rm.cpu(), sd['norm.running_mean']),
                                                                  H.rel_err(rv.cpu(), sd['norm.running_var'])))
        print('   running_mean identical on all ranks: %s' % all(torch.equal(r, rms[0]) for r in rms))


# ------------------------------------------------------------------ 2. gradients vs DDP
def build(mode):
    """mode 'peer': PeerGradSync + FlatGradSync; 'ddp': torch DDP for every parameter."""
    from imaginaire_b200.utils.trainer import get_model_optimizer_and_scheduler, get_trainer
    os.environ['IMX_PEER_GRADS'] = '1' if mode == 'peer' else '0'
    cfg = Config(os.path.join(H.ROOT, 'configs', 'spade_c3_mini.yaml'))
    out = get_model_optimizer_and_scheduler(cfg, seed=0)
    trainer = get_trainer(cfg, *out, (), None)
    trainer.criteria['Perceptual'].model.load_torchvision_state_dict(H.seeded_vgg19_features().state_dict())
    return cfg, trainer


def rank_batch(r):
    g = torch.Generator().manual_seed(100 + r)
    return {'images': torch.rand(2, 3, 256, 256, generator=g) * 2 - 1, 'label': H.onehot_label(200 + r, 2, 35, 256, 256, blocky=8)}


def check_grads():
    if SINGLE:      # one process, the concatenated batch: what every data-parallel variant has to reproduce
        parts = [rank_batch(r) for r in range(SINGLE)]
        batch = {k: torch.cat([p[k] for p in parts]).to(dev) for k in parts[0]}
        eps = H.style_noise(12, 2, 64).repeat(SINGLE, 1).to(dev)
        modes = ('single',)
    else:
        batch = {k: v.to(dev) for k, v in rank_batch(rank).items()}
        eps = H.style_noise(12, 2, 64).to(dev)
        modes = ('ddp', 'peer')
    results = {}
    for mode in modes:
        cfg, trainer = build(mode)
        G = trainer.net_G_module
        G.style_encoder.freeze_random, G.style_encoder.eps = True, eps
        trainer.net_G.train(); trainer.net_D.train()
        grads = {}

        def grab(tag, net, opt):
            def step():
                grads.update({tag + n: p.grad.detach().clone() for n, p in net.named_parameters() if p.grad is not None})
            opt.step = step
        grab('D.', trainer.net_D.module, trainer.opt_D)
        grab('G.', G, trainer.opt_G)
        trainer.dis_update(dict(batch))
        trainer.gen_update(dict(batch))
        torch.cuda.synchronize()
        results[mode] = grads
        for s in list({id(v): v for v in grad_sync._ACTIVE.values()}.values()):
            s.close()
        for f in list(grad_sync._FLAT):
            f.close()
    if SINGLE:
        torch.save({k: v.cpu() for k, v in results['single'].items()}, SINGLE_FILE)
        print('single-GPU gradients of the concatenated batch (%d x 2 images) saved' % SINGLE)
        return
    a, b = results['peer'], results['ddp']
    assert a.keys() == b.keys(), 'different sets of parameters received gradients'
    if os.path.exists(SINGLE_FILE) and rank == 0:
        ref = torch.load(SINGLE_FILE)
        for tag, res in (('peer path', a), ('DistributedDataParallel', b)):
            keys = [k for k in ref if k in res]
            isbig = lambda k: ref[k].dim() == 4 and min(ref[k].shape[0], ref[k].shape[1]) > 8   # noqa: E731
            for name, ks in (('D filters', [k for k in keys if k[0] == 'D' and isbig(k)]),
                             ('D rest', [k for k in keys if k[0] == 'D' and not isbig(k)]),
                             ('G filters', [k for k in keys if k[0] == 'G' and isbig(k)]),
                             ('G rest', [k for k in keys if k[0] == 'G' and not isbig(k)])):
                num = sum(float((res[k].cpu().double() - ref[k].double()).pow(2).sum()) for k in ks)
                den = sum(float(ref[k].double().pow(2).sum()) for k in ks)
                print('   %-24s vs one GPU on the concatenated batch: %-9s rel-L2 %.3e' % (tag, name, (num / max(den, 1e-300)) ** 0.5), flush=True)
    groups = {'D filters (peer-memory reduction)': [], 'D rest (flat all-reduce)': [],
              'G filters (peer-memory reduction)': [], 'G rest (flat all-reduce)': []}
    for k in a:
        big = a[k].dim() == 4 and min(a[k].shape[0], a[k].shape[1]) > 8
        groups[k[0] + (' filters (peer-memory reduction)' if big else ' rest (flat all-reduce)')].append(k)
    say('2. averaged gradients of one iteration (dis_update + gen_update), peer path vs DistributedDataParallel, %d ranks' % world)
    for name, keys in groups.items():
        num = sum(float((a[k].double() - b[k].double()).pow(2).sum()) for k in keys)
        den = sum(float(b[k].double().pow(2).sum()) for k in keys)
        worst = max(keys, key=lambda k: rel_l2(a[k], b[k]) if float(b[k].abs().max()) > 1e-6 else 0.0)
        say('   %-34s %4d tensors  rel-L2 %.3e   worst tensor %.3e (%s)' % (name, len(keys), (num / max(den, 1e-300)) ** 0.5,
                                                                           rel_l2(a[worst], b[worst]), worst))
    # every rank must hold the same averaged gradient
    chk = torch.tensor([sum(float(v.double().sum()) for v in a.values())], device=dev, dtype=torch.float64)
    all_chk = [torch.empty_like(chk) for _ in range(world)]
    dist.all_gather(all_chk, chk)
    say('   gradient checksum per rank (peer path): ' + ' '.join('%.10e' % float(c) for c in all_chk))


if SINGLE:
    check_grads()
else:
    check_syncbn()
    dist.barrier()
    check_grads()
    dist.barrier()
    dist.destroy_process_group()
