"""Tensor-parallel group plumbing: one process per GPU, torch.distributed (NCCL over
NVLink 5 / NVSwitch on the B200 box, gloo in the CPU tests).  Mirrors the call surface the linear
layers use in the reference (python/sglang/srt/distributed/communication_op.py:11-20,
parallel_state.py:398-461, 1047-1200): the all-reduce sits only at the row-parallel boundaries."""
from __future__ import annotations

from typing import Optional

import torch
import torch.distributed as dist

_TP_GROUP: Optional[dist.ProcessGroup] = None
_TP_SIZE = 1
_TP_RANK = 0


def initialize_model_parallel(tensor_model_parallel_size: int = 1,
                              backend: Optional[str] = None) -> None:
    """All ranks of the default group form ONE tensor-parallel group (pp = dp = 1 on this path)."""
    global _TP_GROUP, _TP_SIZE, _TP_RANK
    if tensor_model_parallel_size == 1:
        _TP_GROUP, _TP_SIZE, _TP_RANK = None, 1, 0
        return
    if not dist.is_initialized():
        raise RuntimeError("torch.distributed must be initialised before initialize_model_parallel")
    world = dist.get_world_size()
    if world != tensor_model_parallel_size:
        raise ValueError(f"tp size {tensor_model_parallel_size} != world size {world}")
    _TP_GROUP = dist.group.WORLD
    _TP_SIZE = world
    _TP_RANK = dist.get_rank()


def destroy_model_parallel() -> None:
    global _TP_GROUP, _TP_SIZE, _TP_RANK
    _TP_GROUP, _TP_SIZE, _TP_RANK = None, 1, 0


def get_tensor_model_parallel_world_size() -> int:
    return _TP_SIZE


def get_tensor_model_parallel_rank() -> int:
    return _TP_RANK


def tensor_model_parallel_all_reduce(input_: torch.Tensor) -> torch.Tensor:
    """Sum over the TP group (in place, like the NCCL path of GroupCoordinator.all_reduce)."""
    if _TP_SIZE == 1:
        return input_
    dist.all_reduce(input_, op=dist.ReduceOp.SUM, group=_TP_GROUP)
    return input_


def tensor_model_parallel_all_gather(input_: torch.Tensor, dim: int = -1) -> torch.Tensor:
    if _TP_SIZE == 1:
        return input_
    if dim < 0:
        dim += input_.dim()
    parts = [torch.empty_like(input_) for _ in range(_TP_SIZE)]
    dist.all_gather(parts, input_.contiguous(), group=_TP_GROUP)
    return torch.cat(parts, dim=dim)
