"""TEST INFRASTRUCTURE ONLY (CPU oracle; never imported by the product path).

Restatement of the reference's one-shot custom all-reduce arithmetic
(sgl-kernel/csrc/allreduce/custom_all_reduce.cuh:178-225: `packed_reduce` up-casts every rank's
packed value to fp32, adds them in RANK ORDER starting from rank 0, and down-casts once) and of
the row-parallel linear that feeds it (python/sglang/srt/layers/linear.py:1277-1299: every rank's
GEMM output is rounded to the activation dtype first, bias on rank 0 only, then all-reduced).

Pinned by: the reference's own test of this path compares against NCCL `dist.all_reduce`
(sgl-kernel/tests/test_custom_allreduce.py:144-161, `torch.testing.assert_close`); there are no
stored vectors.  The rank-ordered fp32 sum below is what that kernel computes; NCCL's order is
unspecified, so tests use this oracle bit-exactly and NCCL within tolerance.
"""
from __future__ import annotations

from typing import List

import torch


def all_reduce_oracle(inputs: List[torch.Tensor]) -> torch.Tensor:
    """inputs[r] = rank r's tensor (same shape/dtype) -> the tensor every rank must end up with."""
    acc = torch.zeros(inputs[0].shape, dtype=torch.float32)
    for t in inputs:  # rank order, fp32 accumulation
        acc = acc + t.detach().cpu().to(torch.float32)
    return acc.to(inputs[0].dtype)


def row_parallel_linear_oracle(rank_outputs: List[torch.Tensor]) -> torch.Tensor:
    """rank_outputs[r] = the (already rounded, bias on rank 0) output of rank r's local GEMM."""
    return all_reduce_oracle(rank_outputs)
