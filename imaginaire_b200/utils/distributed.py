"""Process-group helpers.  ref: imaginaire/utils/distributed.py:11-93

One process per GPU, NCCL over NVLink/NVSwitch (the reference's default backend; the
fork's train.py forces gloo, which this does not reproduce).  CPU-only tests pass
backend='gloo'.
"""
import functools
import os

import torch
import torch.distributed as dist


def init_dist(local_rank, backend='nccl', **kwargs):
    if dist.is_available() and not dist.is_initialized():
        if backend == 'nccl':
            torch.cuda.set_device(local_rank)
        dist.init_process_group(backend=backend, init_method='env://', **kwargs)


def get_rank():
    return dist.get_rank() if dist.is_available() and dist.is_initialized() else 0


def get_world_size():
    return dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1


def is_master():
    return get_rank() == 0


def master_only(func):
    @functools.wraps(func)
    def wrapper(*args, **kwargs):
        return func(*args, **kwargs) if is_master() else None
    return wrapper


@master_only
def master_only_print(*args):
    print(*args)


def dist_reduce_tensor(tensor):
    """Mean over ranks, result valid on rank 0."""
    world = get_world_size()
    if world < 2:
        return tensor
    with torch.no_grad():
        dist.reduce(tensor, dst=0)
        if get_rank() == 0:
            tensor /= world
    return tensor


def dist_all_reduce_tensor(tensor):
    """Mean over ranks, on every rank."""
    world = get_world_size()
    if world < 2:
        return tensor
    with torch.no_grad():
        dist.all_reduce(tensor)
        tensor.div_(world)
    return tensor


def dist_all_gather_tensor(tensor):
    world = get_world_size()
    if world < 2:
        return [tensor]
    out = [torch.ones_like(tensor) for _ in range(world)]
    with torch.no_grad():
        dist.all_gather(out, tensor)
    return out
