"""Tensor-parallel group plumbing: one process per GPU, torch.distributed (NCCL over
NVLink 5 / NVSwitch on the B200 box, gloo in the CPU tests).  Mirrors the call surface the linear
layers use in the reference (python/sglang/srt/distributed/communication_op.py:11-20,
parallel_state.py:398-461, 1047-1200): the all-reduce sits only at the row-parallel boundaries."""
from __future__ import annotations

import os
from typing import Optional

import torch
import torch.distributed as dist

_TP_GROUP: Optional[dist.ProcessGroup] = None
_TP_SIZE = 1
_TP_RANK = 0
_TP_CA_COMM = None  # B200AllReduce: NVLink peer-memory all-reduce for decode-sized messages


def initialize_model_parallel(tensor_model_parallel_size: int = 1,
                              backend: Optional[str] = None,
                              use_custom_all_reduce: Optional[bool] = None) -> None:
    """All ranks of the default group form ONE tensor-parallel group (pp = dp = 1 on this path).
    `use_custom_all_reduce` (default: on for NCCL groups, env B200_CUSTOM_AR=0 turns it off)
    builds the peer-memory communicator the way GroupCoordinator.__init__ builds `ca_comm`
    (parallel_state.py:262-268)."""
    global _TP_GROUP, _TP_SIZE, _TP_RANK, _TP_CA_COMM
    if tensor_model_parallel_size == 1:
        _TP_GROUP, _TP_SIZE, _TP_RANK, _TP_CA_COMM = None, 1, 0, None
        return
    if not dist.is_initialized():
        raise RuntimeError("torch.distributed must be initialised before initialize_model_parallel")
    world = dist.get_world_size()
    if world != tensor_model_parallel_size:
        raise ValueError(f"tp size {tensor_model_parallel_size} != world size {world}")
    _TP_GROUP = dist.group.WORLD
    _TP_SIZE = world
    _TP_RANK = dist.get_rank()
    _TP_CA_COMM = None
    if use_custom_all_reduce is None:
        use_custom_all_reduce = (dist.get_backend() == "nccl"
                                 and os.environ.get("B200_CUSTOM_AR", "1") != "0")
    if use_custom_all_reduce:
        from .device_communicators.b200_all_reduce import B200AllReduce

        _TP_CA_COMM = B200AllReduce(_TP_GROUP, torch.device("cuda", torch.cuda.current_device()))


def destroy_model_parallel() -> None:
    global _TP_GROUP, _TP_SIZE, _TP_RANK, _TP_CA_COMM
    if _TP_CA_COMM is not None:
        _TP_CA_COMM.close()
    _TP_GROUP, _TP_SIZE, _TP_RANK, _TP_CA_COMM = None, 1, 0, None


def get_tp_custom_all_reduce():
    """The peer-memory communicator of the TP group (None when disabled / CPU / tp == 1)."""
    if _TP_CA_COMM is not None and not _TP_CA_COMM.disabled:
        return _TP_CA_COMM
    return None


def get_tensor_model_parallel_world_size() -> int:
    return _TP_SIZE


def get_tensor_model_parallel_rank() -> int:
    return _TP_RANK


def tensor_model_parallel_all_reduce(input_: torch.Tensor) -> torch.Tensor:
    """Sum over the TP group: out-of-place through the peer-memory communicator when the message
    qualifies (GroupCoordinator.all_reduce, parallel_state.py:442-449), else in place over NCCL."""
    if _TP_SIZE == 1:
        return input_
    if _TP_CA_COMM is not None and not _TP_CA_COMM.disabled:
        out = _TP_CA_COMM.custom_all_reduce(input_)
        if out is not None:
            return out
    dist.all_reduce(input_, op=dist.ReduceOp.SUM, group=_TP_GROUP)
    return input_


def tensor_model_parallel_all_gather(input_: torch.Tensor, dim: int = -1) -> torch.Tensor:
    if _TP_SIZE == 1:
        return input_
    if dim < 0:
        dim += input_.dim()
    input_ = input_.contiguous()
    if input_.is_cuda:
        # one flat collective into [tp, ...] then a single reshuffle (GroupCoordinator.all_gather,
        # parallel_state.py:463-500) — capturable in a CUDA graph
        out = torch.empty((_TP_SIZE,) + tuple(input_.shape), dtype=input_.dtype,
                          device=input_.device)
        dist.all_gather_into_tensor(out, input_, group=_TP_GROUP)
        out = out.movedim(0, dim)
        shape = list(input_.shape)
        shape[dim] *= _TP_SIZE
        return out.reshape(shape)
    parts = [torch.empty_like(input_) for _ in range(_TP_SIZE)]
    dist.all_gather(parts, input_, group=_TP_GROUP)
    return torch.cat(parts, dim=dim)
