"""Field-compatible stand-in for the reference's ForwardBatch / ForwardMode.

Only the fields the attention backend consumes are kept (reference
python/sglang/srt/model_executor/forward_batch_info.py:54-118 for the mode predicates and
:139-259 for the batch fields).  In a real deployment the reference's own classes are passed in;
these exist so that tests, bench.py and smoke() can drive the backend without importing sglang.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from enum import IntEnum, auto
from typing import Any, List, Optional

import torch


class ForwardMode(IntEnum):
    EXTEND = auto()         # prefill of new tokens after an already cached prefix
    DECODE = auto()         # one new token per request
    MIXED = auto()          # chunked prefill: extends and 1-token "extends" of decodes
    IDLE = auto()
    TARGET_VERIFY = auto()
    DRAFT_EXTEND = auto()
    DUMMY_FIRST = auto()

    def is_extend(self) -> bool:
        return self in (ForwardMode.EXTEND, ForwardMode.MIXED, ForwardMode.DRAFT_EXTEND,
                        ForwardMode.TARGET_VERIFY)

    def is_prefill(self) -> bool:
        return self.is_extend()

    def is_decode(self) -> bool:
        return self == ForwardMode.DECODE

    def is_mixed(self) -> bool:
        return self == ForwardMode.MIXED

    def is_idle(self) -> bool:
        return self == ForwardMode.IDLE

    def is_cuda_graph(self) -> bool:
        return self in (ForwardMode.DECODE, ForwardMode.TARGET_VERIFY, ForwardMode.IDLE)

    def is_decode_or_idle(self) -> bool:
        return self in (ForwardMode.DECODE, ForwardMode.IDLE)


@dataclass
class ForwardBatch:
    forward_mode: ForwardMode
    batch_size: int
    req_pool_indices: torch.Tensor            # int64 [bs]
    seq_lens: torch.Tensor                    # int64 [bs], total length incl. new tokens
    out_cache_loc: torch.Tensor               # int64 [num_tokens] KV slot of every new token
    seq_lens_sum: int = 0
    seq_lens_cpu: Optional[torch.Tensor] = None
    positions: Optional[torch.Tensor] = None
    # extend only
    extend_seq_lens: Optional[torch.Tensor] = None      # int32 [bs]
    extend_prefix_lens: Optional[torch.Tensor] = None   # int32 [bs]
    extend_start_loc: Optional[torch.Tensor] = None     # int32 [bs]
    extend_seq_lens_cpu: Optional[List[int]] = None
    extend_prefix_lens_cpu: Optional[List[int]] = None
    extend_num_tokens: Optional[int] = None
    # pools / backend
    req_to_token_pool: Any = None
    token_to_kv_pool: Any = None
    attn_backend: Any = None
    spec_info: Any = None
    encoder_lens: Optional[torch.Tensor] = None
    extra: dict = field(default_factory=dict)
