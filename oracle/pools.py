"""CPU restatement of the reference pools (python/sglang/srt/mem_cache/memory_pool.py).

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py)."""
from __future__ import annotations

import torch


class OracleReqToTokenPool:
    # memory_pool.py:54-99 — int32 [size, max_context_len] zero-initialised table
    def __init__(self, size: int, max_context_len: int):
        self.size = size
        self.max_context_len = max_context_len
        self.req_to_token = torch.zeros((size, max_context_len), dtype=torch.int32)

    def write(self, indices, values):
        self.req_to_token[indices] = values


class OracleMHATokenToKVPool:
    # memory_pool.py:238-412 — per layer [size + page_size, Hkv, D]; slot 0 is the dummy row
    def __init__(self, size, page_size, dtype, head_num, head_dim, layer_num):
        self.size, self.page_size, self.dtype = size, page_size, dtype
        self.head_num, self.head_dim, self.layer_num = head_num, head_dim, layer_num
        self.start_layer = 0
        shape = (size + page_size, head_num, head_dim)
        # fp8 caches are stored as bytes (memory_pool.py:119-123: index_put has no fp8 kernel)
        self.store_dtype = (torch.uint8 if dtype in (torch.float8_e5m2, torch.float8_e4m3fn)
                            else dtype)
        self.k_buffer = [torch.zeros(shape, dtype=self.store_dtype) for _ in range(layer_num)]
        self.v_buffer = [torch.zeros(shape, dtype=self.store_dtype) for _ in range(layer_num)]

    def get_key_buffer(self, layer_id):  # memory_pool.py:359-365
        t = self.k_buffer[layer_id - self.start_layer]
        return t.view(self.dtype) if self.store_dtype != self.dtype else t

    def get_value_buffer(self, layer_id):  # memory_pool.py:367-373
        t = self.v_buffer[layer_id - self.start_layer]
        return t.view(self.dtype) if self.store_dtype != self.dtype else t

    def set_kv_buffer(self, layer, loc, cache_k, cache_v, k_scale=None, v_scale=None):
        # memory_pool.py:378-412 (bf16/f16 cache: plain index_put of the rows; fp8 cache: in-place
        # divide by the scale, cast, store the bytes)
        if cache_k.dtype != self.dtype:
            if k_scale is not None:
                cache_k.div_(k_scale)
            if v_scale is not None:
                cache_v.div_(v_scale)
            cache_k, cache_v = cache_k.to(self.dtype), cache_v.to(self.dtype)
        if self.store_dtype != self.dtype:  # memory_pool.py:398-400
            cache_k, cache_v = cache_k.view(self.store_dtype), cache_v.view(self.store_dtype)
        self.k_buffer[layer.layer_id - self.start_layer][loc] = cache_k
        self.v_buffer[layer.layer_id - self.start_layer][loc] = cache_v


def kv_indices_oracle(req_to_token, req_pool_indices, lens, kv_start=None):
    """Flat gather of each request's token slots: the python-loop definition used as the
    reference in test/srt/test_create_kvindices.py:40-49 (kernel: layers/attention/utils.py:5-41).
    Returns (kv_indptr int32 [bs+1], kv_indices int32 [sum lens])."""
    rows = []
    for b in range(len(req_pool_indices)):
        s = 0 if kv_start is None else int(kv_start[b])
        rows.append(req_to_token[int(req_pool_indices[b]), s:s + int(lens[b])])
    indptr = torch.zeros(len(rows) + 1, dtype=torch.int32)
    indptr[1:] = torch.cumsum(torch.tensor([r.numel() for r in rows], dtype=torch.int64), 0)
    flat = torch.cat(rows).contiguous() if rows else torch.empty(0, dtype=torch.int32)
    return indptr, flat
