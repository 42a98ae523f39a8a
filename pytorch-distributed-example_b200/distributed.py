"""Stand-in for `torch.distributed` as the reference scripts use it.

    from torch import distributed, nn              ->  from torch import nn
                                                       from pytorch_distributed_example_b200 import distributed
    import torch.distributed as dist               ->  from pytorch_distributed_example_b200 import distributed as dist

Names the reference touches and what they do here:
  init_process_group  mnist/main.py:199-203, toy/main.py:29-33  same rendezvous through torch (its
                      group carries the IPC handles), plus the device selection the reference
                      forgets (every rank would sit on GPU 0)
  get_world_size      mnist/main.py:123
  all_reduce          mnist/main.py:126, toy/main.py:20         SUM of fp32/int32 through the fused
                      kernel (one-tensor bucket); other ops/dtypes are rejected, not routed elsewhere
  reduce_op/ReduceOp  mnist/main.py:126, toy/main.py:20
  new_group           toy/main.py:16                            the all-ranks group is a cheap token
  broadcast           (absent from the reference, which relies on identical seeds it never sets,
                      mnist/main.py:185)                        rank `src`'s tensor into every rank's, over the
                      same peer mappings
Everything else resolves to `torch.distributed` unchanged.
"""
import os

import torch
import torch.distributed as _td

from . import _lib
from ._lib import GradAvgError

ReduceOp = _td.ReduceOp


class _ReduceOpAlias:
    """`distributed.reduce_op.SUM` (deprecated spelling the reference uses) without the warning."""
    SUM, PRODUCT, MIN, MAX = _td.ReduceOp.SUM, _td.ReduceOp.PRODUCT, _td.ReduceOp.MIN, _td.ReduceOp.MAX


reduce_op = _ReduceOpAlias()


class WorldToken:
    """What `new_group(all ranks)` returns: the reference makes one per step (toy/main.py:15-16),
    which in torch is a store-backed collective that leaks a group each time."""
    def __init__(self, world_size):
        self.world_size = world_size


def init_process_group(backend=None, init_method=None, rank=-1, world_size=-1, **kwargs):
    if torch.cuda.is_available():
        local = int(os.environ.get("LOCAL_RANK", rank if rank is not None and rank >= 0 else 0))
        torch.cuda.set_device(local % torch.cuda.device_count())        # F5
    kw = dict(kwargs)
    if init_method is not None:
        kw["init_method"] = init_method
    if rank is not None and rank >= 0:
        kw["rank"] = rank
    if world_size is not None and world_size > 0:
        kw["world_size"] = world_size
    return _td.init_process_group(backend=backend, **kw)


def new_group(ranks=None, **kwargs):
    ws = _td.get_world_size()
    if ranks is None or sorted(ranks) == list(range(ws)):
        return WorldToken(ws)
    return _td.new_group(ranks=ranks, **kwargs)


def all_reduce(tensor, op=_td.ReduceOp.SUM, group=None, async_op=False):
    """In-place SUM across ranks, blocking like the call it replaces (async_op unsupported)."""
    from .averager import default_averager
    if async_op:
        raise GradAvgError(_lib.GA_EINVAL, "async_op=True is not supported on this path")
    if op != _td.ReduceOp.SUM:
        raise GradAvgError(_lib.GA_EINVAL, "only ReduceOp.SUM is implemented (the reference uses nothing else)")
    if group is not None and not isinstance(group, WorldToken):
        raise GradAvgError(_lib.GA_EINVAL, "only the default / all-ranks group is supported")
    if tensor.dtype not in (torch.float32, torch.int32):
        raise GradAvgError(_lib.GA_EINVAL, "dtype %s: only float32 and int32 are implemented" % tensor.dtype)
    avg = default_averager()
    if tensor.is_cuda:
        avg.all_reduce_([[tensor]])
        return None
    # Host tensor (toy/main.py:19 builds a CPU IntTensor): stage through the device, no CPU maths.
    staged = tensor.to(avg.device, non_blocking=False).contiguous()
    avg.all_reduce_([[staged]])
    tensor.copy_(staged)            # synchronises: the call is blocking, as on gloo
    return None


def broadcast(tensor, src=0, group=None, async_op=False):
    """In-place broadcast of rank `src`'s tensor (the initial-parameter sync the reference lacks)."""
    from .averager import default_averager
    if async_op:
        raise GradAvgError(_lib.GA_EINVAL, "async_op=True is not supported on this path")
    if group is not None and not isinstance(group, WorldToken):
        raise GradAvgError(_lib.GA_EINVAL, "only the default / all-ranks group is supported")
    avg = default_averager()
    if tensor.is_cuda:
        avg.broadcast_([[tensor]], root=src)
        return None
    staged = tensor.to(avg.device).contiguous()
    avg.broadcast_([[staged]], root=src)
    tensor.copy_(staged)
    return None


def __getattr__(name):
    return getattr(_td, name)
