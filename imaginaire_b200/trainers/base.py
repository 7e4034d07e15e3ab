"""Trainer base class with the reference's interface.
ref: imaginaire/trainers/base.py:27-829 (BaseTrainer, _save_checkpoint)

What the training loop of train.py:63-84 calls is here with the same names, arguments and order of effects:
``start_of_epoch / start_of_iteration / dis_update / gen_update / end_of_iteration / end_of_epoch``,
``save_checkpoint / load_checkpoint`` (same checkpoint keys and ``latest_checkpoint.txt``), gradient-norm clipping
(``gen_opt.clip_grad_norm``, base.py:618-620) and the learning-rate schedulers stepped in iteration or epoch mode.
apex.amp's loss scaling has no counterpart: the kernels compute in bf16 storage with fp32 accumulation and need no
scale.  Tensorboard meters, image dumps and FID are outside the hot path (out of scope) and are
no-ops.
"""
import os
import time

import torch
from torch import nn

from .. import grad_sync
from ..utils.distributed import is_master, master_only
from ..utils.distributed import master_only_print as print
from ..utils.model_average import requires_grad


def to_cuda(data):
    """Moves every tensor of a (nested) batch to the current CUDA device.  ref: imaginaire/utils/misc.py to_cuda"""
    if isinstance(data, torch.Tensor):
        return data if data.is_cuda else data.cuda(non_blocking=True)
    if isinstance(data, dict):
        return {k: to_cuda(v) for k, v in data.items()}
    if isinstance(data, (list, tuple)):
        return type(data)(to_cuda(v) for v in data)
    return data


class BaseTrainer(object):
    def __init__(self, cfg, net_G, net_D, opt_G, opt_D, sch_G, sch_D, train_data_loader, val_data_loader):
        super().__init__()
        self.cfg = cfg
        self.net_G = net_G
        # two wrappers (data parallel + model average) or one (base.py:58-63)
        self.net_G_module = self.net_G.module.module if cfg.trainer.model_average else self.net_G.module
        self.val_data_loader = val_data_loader
        self.is_inference = train_data_loader is None
        self.net_D = net_D
        self.opt_G, self.opt_D = opt_G, opt_D
        self.sch_G, self.sch_D = sch_G, sch_D
        self.train_data_loader = train_data_loader

        self.criteria = nn.ModuleDict()
        self.weights = dict()
        self.losses = dict(gen_update=dict(), dis_update=dict())
        self.gen_losses = self.losses['gen_update']
        self.dis_losses = self.losses['dis_update']
        self._init_loss(cfg)
        dev = next(self.net_G.parameters()).device
        for loss_name in self.weights:
            if loss_name in self.criteria and self.criteria[loss_name] is not None:
                self.criteria[loss_name].to(dev)
        if self.is_inference:
            return
        self.current_iteration = 0
        self.current_epoch = 0
        self.start_iteration_time = None
        self.start_epoch_time = None
        self.elapsed_iteration_time = 0
        self.time_iteration = -1
        self.time_epoch = -1

    # ------------------------------------------------------------------ hooks of the child class
    def _init_loss(self, cfg):
        raise NotImplementedError

    def gen_forward(self, data):
        raise NotImplementedError

    def dis_forward(self, data):
        raise NotImplementedError

    def _start_of_epoch(self, current_epoch):
        pass

    def _start_of_iteration(self, data, current_iteration):
        return data

    def _end_of_iteration(self, data, current_epoch, current_iteration):
        pass

    def _end_of_epoch(self, data, current_epoch, current_iteration):
        pass

    def _pre_save_checkpoint(self):
        pass

    # ------------------------------------------------------------------ checkpoints (base.py:199-265, 795-829)
    def save_checkpoint(self, current_epoch, current_iteration):
        self._pre_save_checkpoint()
        return _save_checkpoint(self.cfg, self.net_G, self.net_D, self.opt_G, self.opt_D, self.sch_G, self.sch_D,
                                current_epoch, current_iteration)

    def load_checkpoint(self, cfg, checkpoint_path, resume=None):
        if os.path.exists(checkpoint_path):
            if resume is None:
                resume = False
        elif os.path.exists(os.path.join(cfg.logdir, 'latest_checkpoint.txt')):
            with open(os.path.join(cfg.logdir, 'latest_checkpoint.txt'), 'r') as f:
                line = f.read().splitlines()
            checkpoint_path = os.path.join(cfg.logdir, line[0].split(' ')[-1])
            if resume is None:
                resume = True
        else:
            print('No checkpoint found.')
            return 0, 0
        checkpoint = torch.load(checkpoint_path, map_location=lambda storage, loc: storage, weights_only=False)
        current_epoch = current_iteration = 0
        if resume:
            self.net_G.load_state_dict(checkpoint['net_G'])
            if not self.is_inference:
                self.net_D.load_state_dict(checkpoint['net_D'])
                if 'opt_G' in checkpoint:
                    self.opt_G.load_state_dict(checkpoint['opt_G'])
                    self.opt_D.load_state_dict(checkpoint['opt_D'])
                    self.sch_G.load_state_dict(checkpoint['sch_G'])
                    self.sch_D.load_state_dict(checkpoint['sch_D'])
                    current_epoch = checkpoint['current_epoch']
                    current_iteration = checkpoint['current_iteration']
                    print('Load from: {}'.format(checkpoint_path))
                else:
                    print('Load network weights only.')
        else:
            self.net_G.load_state_dict(checkpoint['net_G'])
            print('Load generator weights only.')
        print('Done with loading the checkpoint.')
        return current_epoch, current_iteration

    # ------------------------------------------------------------------ loop hooks (base.py:267-401)
    def start_of_epoch(self, current_epoch):
        self._start_of_epoch(current_epoch)
        self.current_epoch = current_epoch
        self.start_epoch_time = time.time()

    def start_of_iteration(self, data, current_iteration):
        data = self._start_of_iteration(data, current_iteration)
        data = to_cuda(data)
        self.current_iteration = current_iteration
        if not self.is_inference:
            self.net_D.train()
        self.net_G.train()
        self.start_iteration_time = time.time()
        return data

    def end_of_iteration(self, data, current_epoch, current_iteration):
        self.current_iteration = current_iteration
        self.current_epoch = current_epoch
        if self.cfg.gen_opt.lr_policy.iteration_mode:
            self.sch_G.step()
        if self.cfg.dis_opt.lr_policy.iteration_mode:
            self.sch_D.step()
        self.elapsed_iteration_time += time.time() - self.start_iteration_time
        if current_iteration % self.cfg.logging_iter == 0:
            self.time_iteration = self.elapsed_iteration_time / self.cfg.logging_iter
            print('Iteration: {}, average iter time: {:6f}.'.format(current_iteration, self.time_iteration))
            self.elapsed_iteration_time = 0
        self._end_of_iteration(data, current_epoch, current_iteration)
        if current_iteration >= getattr(self.cfg, 'snapshot_save_start_iter', 0) and \
                current_iteration % self.cfg.snapshot_save_iter == 0:
            self.save_checkpoint(current_epoch, current_iteration)

    def end_of_epoch(self, data, current_epoch, current_iteration):
        self.current_iteration = current_iteration
        self.current_epoch = current_epoch
        if not self.cfg.gen_opt.lr_policy.iteration_mode:
            self.sch_G.step()
        if not self.cfg.dis_opt.lr_policy.iteration_mode:
            self.sch_D.step()
        self.time_epoch = time.time() - self.start_epoch_time
        print('Epoch: {}, total time: {:6f}.'.format(current_epoch, self.time_epoch))
        self._end_of_epoch(data, current_epoch, current_iteration)
        if current_epoch >= getattr(self.cfg, 'snapshot_save_start_epoch', 0) and \
                current_epoch % self.cfg.snapshot_save_epoch == 0:
            self.save_checkpoint(current_epoch, current_iteration)

    # ------------------------------------------------------------------ the two updates (base.py:594-666)
    def gen_update(self, data):
        self.opt_G.zero_grad(set_to_none=True)
        requires_grad(self.net_G_module, True)
        requires_grad(self.net_D, False)
        total_loss = self.gen_forward(data)
        if total_loss is None:
            return
        total_loss.backward()
        grad_sync.finish_all()      # no-op unless the peer-memory gradient path is active
        if hasattr(self.cfg.gen_opt, 'clip_grad_norm'):
            nn.utils.clip_grad_norm_([p for g in self.opt_G.param_groups for p in g['params']],
                                     self.cfg.gen_opt.clip_grad_norm)
        self.opt_G.step()
        if self.cfg.trainer.model_average:
            self.net_G.module.update_average()
        self._detach_losses()

    def dis_update(self, data):
        self.opt_D.zero_grad(set_to_none=True)
        requires_grad(self.net_G_module, False)
        requires_grad(self.net_D, True)
        total_loss = self.dis_forward(data)
        if total_loss is None:
            return
        total_loss.backward()
        grad_sync.finish_all()
        self.opt_D.step()
        self._detach_losses()

    def _get_outputs(self, net_D_output, real=True):
        """Discriminator outputs to score; differences real - fake / fake - real in relativistic mode
        (base.py:498-536)."""
        def difference(a, b):
            return [difference(x, y) if isinstance(x, (list, tuple)) else x - y for x, y in zip(a, b)]
        first, second = ('real_outputs', 'fake_outputs') if real else ('fake_outputs', 'real_outputs')
        if self.cfg.trainer.gan_relativistic:
            return difference(net_D_output[first], net_D_output[second])
        return net_D_output[first]

    def _detach_losses(self):
        """Loss tensors keep no graph after the update (base.py:_detach_losses)."""
        for losses in self.losses.values():
            for name, value in losses.items():
                losses[name] = value.detach()


@master_only
def _save_checkpoint(cfg, net_G, net_D, opt_G, opt_D, sch_G, sch_D, current_epoch, current_iteration):
    """Same file name pattern, dictionary keys and latest_checkpoint.txt as the reference (base.py:795-829): the
    state dicts carry the wrappers' ``module.`` / ``module.module.`` / ``module.averaged_model.`` prefixes."""
    latest_checkpoint_path = 'epoch_{:05}_iteration_{:09}_checkpoint.pt'.format(current_epoch, current_iteration)
    os.makedirs(cfg.logdir, exist_ok=True)
    save_path = os.path.join(cfg.logdir, latest_checkpoint_path)
    torch.save({'net_G': net_G.state_dict(), 'net_D': net_D.state_dict(),
                'opt_G': opt_G.state_dict(), 'opt_D': opt_D.state_dict(),
                'sch_G': sch_G.state_dict(), 'sch_D': sch_D.state_dict(),
                'current_epoch': current_epoch, 'current_iteration': current_iteration}, save_path)
    with open(os.path.join(cfg.logdir, 'latest_checkpoint.txt'), 'wt') as f:
        f.write('latest_checkpoint: %s' % latest_checkpoint_path)
    print('Save checkpoint to {}'.format(save_path))
    return save_path


__all__ = ['BaseTrainer', 'to_cuda', 'is_master']
