"""AttentionBackend interface (mirror of the reference ABC,
python/sglang/srt/layers/attention/base_attn_backend.py:14-115): same method names, argument
order and dispatch rule, so a backend written against it drops into ModelRunner unchanged."""
from __future__ import annotations

from abc import ABC, abstractmethod


class AttentionBackend(ABC):
    @abstractmethod
    def init_forward_metadata(self, forward_batch):
        """Prepare per-step metadata; called once per forward before layer 0."""

    def init_cuda_graph_state(self, max_bs: int):
        raise NotImplementedError()

    def init_forward_metadata_capture_cuda_graph(self, bs, num_tokens, req_pool_indices, seq_lens,
                                                 encoder_lens, forward_mode, spec_info):
        raise NotImplementedError()

    def init_forward_metadata_replay_cuda_graph(self, bs, req_pool_indices, seq_lens, seq_lens_sum,
                                                encoder_lens, forward_mode, spec_info,
                                                seq_lens_cpu):
        raise NotImplementedError()

    def get_cuda_graph_seq_len_fill_value(self):
        raise NotImplementedError()

    def forward(self, q, k, v, layer, forward_batch, save_kv_cache: bool = True, **kwargs):
        # decode mode -> forward_decode, every other mode (extend / mixed / verify) -> extend
        fn = self.forward_decode if forward_batch.forward_mode.is_decode() else self.forward_extend
        return fn(q, k, v, layer, forward_batch, save_kv_cache=save_kv_cache, **kwargs)

    def forward_decode(self, q, k, v, layer, forward_batch, save_kv_cache: bool = True):
        raise NotImplementedError()

    def forward_extend(self, q, k, v, layer, forward_batch, save_kv_cache: bool = True):
        raise NotImplementedError()

    def support_triton(self):
        return True
