"""Builds networks, optimisers and the data-parallel wrapping.
ref: imaginaire/utils/trainer.py:24-37 (seeds), :69-125 (get_model_optimizer_and_scheduler),
     :128-216 (wrap / DDP), :219-306 (schedulers, optimisers), :40-66 (get_trainer)

apex is replaced as SURVEY.md Appendix C prescribes: mixed precision is the kernels' own
bf16-storage / fp32-accumulate arithmetic (no loss scaling), apex FusedAdam becomes
imaginaire_b200.optim.FusedAdam (csrc/optim.cu), apex DDP becomes torch DDP over NCCL.
"""
import importlib
import os
import random

import numpy as np
import torch
from torch.nn.parallel import DistributedDataParallel
from torch.optim import lr_scheduler

from .distributed import get_rank, get_world_size
from .init_weight import weights_init
from .model_average import ModelAverage
from ..optim import FusedAdam


def set_random_seed(seed, by_rank=False):
    if by_rank:
        seed += get_rank()
    random.seed(seed)
    np.random.seed(seed)
    torch.manual_seed(seed)
    torch.cuda.manual_seed_all(seed)


def get_optimizer(cfg_opt, net):
    """ref: utils/trainer.py:261-276.  Every parameter of the network goes to the optimizer (also the ones that
    are frozen at construction: the trainer flips requires_grad per phase, trainers/base.py:602-603)."""
    if hasattr(net, 'get_param_groups'):
        params = net.get_param_groups(cfg_opt)
    elif isinstance(net, torch.nn.Module):
        params = net.parameters()
    else:
        params = net        # an iterable of parameters / parameter groups
    return get_optimizer_for_params(cfg_opt, params)


def get_optimizer_for_params(cfg_opt, params):
    """ref: utils/trainer.py:279-306.  'adam' is the library's multi-tensor kernel (the reference's apex FusedAdam /
    torch Adam: same update rule and state layout); like every other product path it refuses CPU parameters at
    step().  The other optimizers of the reference (madam, fromage, rmsprop, sgd) are outside the SPADE path."""
    if cfg_opt.type != 'adam':
        raise NotImplementedError('Optimizer {} is not wired up for the SPADE step.'.format(cfg_opt.type))
    return FusedAdam(list(params), lr=cfg_opt.lr, eps=getattr(cfg_opt, 'eps', 1e-8),
                     betas=(cfg_opt.adam_beta1, cfg_opt.adam_beta2))


def get_scheduler(cfg_opt, opt):
    """ref: utils/trainer.py:219-240: 'step' -> StepLR(step_size, gamma), 'constant' -> LambdaLR(1)."""
    policy = cfg_opt.lr_policy
    if policy.type == 'step':
        return lr_scheduler.StepLR(opt, step_size=policy.step_size, gamma=policy.gamma)
    if policy.type == 'constant':
        return lr_scheduler.LambdaLR(opt, lambda x: 1)
    raise NotImplementedError('Learning rate policy {} not implemented.'.format(policy.type))


def get_trainer(cfg, net_G, net_D=None, opt_G=None, opt_D=None, sch_G=None, sch_D=None, train_data_loader=None,
                val_data_loader=None):
    """ref: utils/trainer.py:40-66: the Trainer class of the module that ``cfg.trainer.type`` names (resolved through
    the ``imaginaire`` alias package, so the reference's config files work unchanged)."""
    trainer_lib = importlib.import_module(_resolve(cfg.trainer.type))
    return trainer_lib.Trainer(cfg, net_G, net_D, opt_G, opt_D, sch_G, sch_D, train_data_loader, val_data_loader)


def _resolve(module_name):
    """'imaginaire.x.y' -> 'imaginaire_b200.x.y' (the same module object the alias package would return)."""
    if module_name == 'imaginaire' or module_name.startswith('imaginaire.'):
        return 'imaginaire_b200' + module_name[len('imaginaire'):]
    return module_name


def get_model_optimizer_and_scheduler(cfg, seed=0, device='cuda'):
    """ref: utils/trainer.py:69-125.  Returns (net_G, net_D, opt_G, opt_D, sch_G, sch_D).  Networks are built by the
    modules ``cfg.gen.type`` / ``cfg.dis.type`` name and initialised with the same seed on every rank, then the seed
    is offset by rank like the reference."""
    set_random_seed(seed, by_rank=False)
    lib_G = importlib.import_module(_resolve(cfg.gen.type))
    lib_D = importlib.import_module(_resolve(cfg.dis.type))
    net_G = lib_G.Generator(cfg.gen, cfg.data)
    net_D = lib_D.Discriminator(cfg.dis, cfg.data)
    init = getattr(cfg.trainer, 'init', None)
    if init is not None:
        fn = weights_init(init.type, init.gain, getattr(init, 'bias', None))
        net_G.apply(fn)
        net_D.apply(fn)
    net_G, net_D = net_G.to(device), net_D.to(device)
    set_random_seed(seed, by_rank=True)
    opt_G = get_optimizer(cfg.gen_opt, net_G)
    opt_D = get_optimizer(cfg.dis_opt, net_D)
    net_G, net_D, opt_G, opt_D = wrap_model_and_optimizer(cfg, net_G, net_D, opt_G, opt_D)
    return net_G, net_D, opt_G, opt_D, get_scheduler(cfg.gen_opt, opt_G), get_scheduler(cfg.dis_opt, opt_D)


def wrap_model_and_optimizer(cfg, net_G, net_D, opt_G, opt_D):
    if cfg.trainer.model_average:
        net_G = ModelAverage(net_G, cfg.trainer.model_average_beta,
                             cfg.trainer.model_average_start_iteration,
                             getattr(cfg.trainer, 'model_average_remove_sn', True))
    find_unused = getattr(cfg.trainer, 'ddp_find_unused_parameters', None)
    # the discriminator is applied to real and fake images and its FPSE heads to three pyramid levels
    # each: up to 6 applications of one filter per backward, each with its own gradient slot
    return _wrap_model(net_G, find_unused, 1), _wrap_model(net_D, find_unused, 8), opt_G, opt_D


class WrappedModel(torch.nn.Module):
    """Single-process stand-in for DDP so that ``.module`` exists either way (trainer.py:168-190)."""

    def __init__(self, module):
        super().__init__()
        self.module = module

    def forward(self, *args, **kwargs):
        return self.module(*args, **kwargs)


def _wrap_model(model, find_unused=None, max_uses=1):
    """find_unused: None = the reference's setting (True, utils/trainer.py:203-214) unless the
    environment variable IMX_DDP_FIND_UNUSED=0 says that every trainable parameter takes part in every
    backward (true for the SPADE trainer here, which calls a frozen network through .module, not
    through the DDP wrapper): the per-iteration autograd-graph walk is then skipped."""
    if get_world_size() > 1:
        if find_unused is None:
            find_unused = os.environ.get('IMX_DDP_FIND_UNUSED', '1') != '0'
        if os.environ.get('IMX_PEER_GRADS', '1') != '0' and get_world_size() <= 8:
            # filter gradients are averaged over NVLink peer memory inside the weight-gradient reduction
            # (grad_sync.py); DDP keeps the small rest.  If symmetric memory cannot be set up on any rank,
            # every rank falls back to plain DDP.
            import torch.distributed as dist
            from .. import grad_sync
            filters = grad_sync.eligible_filters(model)
            sync, ok = None, torch.ones(1, device='cuda')
            flat_grads = os.environ.get('IMX_FLAT_GRADS', '1') != '0'
            carried = {id(p) for _, p in filters}
            if filters:
                try:
                    sync = grad_sync.PeerGradSync(
                        [p for _, p in filters], max_uses=max_uses,
                        extra_params=[p for p in model.parameters() if id(p) not in carried] if flat_grads else None,
                        workspace_bytes=int(float(os.environ.get('IMX_PEER_GRADS_WS_GB', '16' if max_uses == 1 else '8')) * (1 << 30)))
                except Exception as e:     # noqa: BLE001 - any failure means "use NCCL for everything"
                    print('peer-memory gradient path unavailable (%s); using DDP for all parameters' % repr(e)[:200])
                    ok.zero_()
                dist.all_reduce(ok, op=dist.ReduceOp.MIN)
                if float(ok) > 0:
                    if flat_grads:
                        # everything else rides the same reduce-scatter / all-gather as one more segment of the arena
                        # (PeerGradSync extra_params) instead of a DDP wrapper: no autograd hooks / buckets /
                        # unused-parameter search, no NCCL call, and the iteration stays capturable as a CUDA graph
                        grad_sync.broadcast_state(model)
                        return WrappedModel(model)
                    DistributedDataParallel._set_params_and_buffers_to_ignore_for_model(model, [n for n, _ in filters])
                elif sync is not None:
                    sync.close()
        ddp = DistributedDataParallel(model, device_ids=[torch.cuda.current_device()],
                                      output_device=torch.cuda.current_device(),
                                      find_unused_parameters=bool(find_unused), broadcast_buffers=False,
                                      gradient_as_bucket_view=True, bucket_cap_mb=int(os.environ.get('IMX_DDP_BUCKET_MB', '256')))
        if os.environ.get('IMX_DDP_BF16', '0') == '1':
            from torch.distributed.algorithms.ddp_comm_hooks import default_hooks
            ddp.register_comm_hook(None, default_hooks.bf16_compress_hook)
        return ddp
    return WrappedModel(model)
