"""KV / request pools with the reference's layout (the layout is the drop-in contract).

  ReqToTokenPool.req_to_token : int32 [size, max_context_len], row = request slot, column = token
                                position, value = KV slot id          (memory_pool.py:54-99)
  MHATokenToKVPool.k_buffer[l], v_buffer[l] : [size + page_size, Hkv, D], slot 0 = dummy row for
                                padded tokens                          (memory_pool.py:238-412)
  TokenToKVPoolAllocator      : free slots 1..size                      (memory_pool.py:169-235)

`set_kv_buffer` is the cache-location write of the hot path and goes through the C-ABI
(b200_kv_write): a bit-exact row copy.  These classes only run on CUDA tensors.
"""
from __future__ import annotations

from typing import List, Optional, Union

import torch

from .. import _lib


class ReqToTokenPool:
    def __init__(self, size: int, max_context_len: int, device: str,
                 enable_memory_saver: bool = False):
        self.size = size
        self.max_context_len = max_context_len
        self.device = device
        self.req_to_token = torch.zeros((size, max_context_len), dtype=torch.int32, device=device)
        self.free_slots = list(range(size))

    def write(self, indices, values):
        self.req_to_token[indices] = values

    def available_size(self) -> int:
        return len(self.free_slots)

    def alloc(self, need_size: int) -> Optional[List[int]]:
        if need_size > len(self.free_slots):
            return None
        taken, self.free_slots = self.free_slots[:need_size], self.free_slots[need_size:]
        return taken

    def free(self, free_index: Union[int, List[int]]):
        if isinstance(free_index, int):
            self.free_slots.append(free_index)
        else:
            self.free_slots.extend(free_index)

    def clear(self):
        self.free_slots = list(range(self.size))


class TokenToKVPoolAllocator:
    """Hands out KV slot ids (int64).  Slot 0 is never handed out: it is the dummy row."""

    def __init__(self, size: int, dtype: torch.dtype, device: str, kvcache=None):
        self.size = size
        self.dtype = dtype
        self.device = device
        self.page_size = 1
        self._kvcache = kvcache
        self.clear()

    def available_size(self) -> int:
        return len(self.free_slots)

    def get_kvcache(self):
        return self._kvcache

    def alloc(self, need_size: int) -> Optional[torch.Tensor]:
        if need_size > len(self.free_slots):
            return None
        out = self.free_slots[:need_size]
        self.free_slots = self.free_slots[need_size:]
        return out

    def free(self, free_index: torch.Tensor):
        if free_index.numel():
            self.free_slots = torch.cat((self.free_slots, free_index))

    def clear(self):
        self.free_slots = torch.arange(1, self.size + 1, dtype=torch.int64, device=self.device)


class MHATokenToKVPool:
    def __init__(self, size: int, page_size: int, dtype: torch.dtype, head_num: int, head_dim: int,
                 layer_num: int, device: str, enable_memory_saver: bool = False,
                 start_layer: Optional[int] = None, end_layer: Optional[int] = None):
        self.size = size
        self.page_size = page_size
        self.dtype = dtype
        self.device = device
        # fp8 caches are stored as bytes, like the reference (index_put has no fp8 kernel there)
        self.store_dtype = torch.uint8 if dtype in (torch.float8_e5m2, torch.float8_e4m3fn) else dtype
        self.layer_num = layer_num
        self.start_layer = start_layer or 0
        self.end_layer = end_layer or layer_num - 1
        self.head_num = head_num
        self.head_dim = head_dim
        self.layer_transfer_counter = None
        shape = (size + page_size, head_num, head_dim)
        self.k_buffer = [torch.zeros(shape, dtype=self.store_dtype, device=device)
                         for _ in range(layer_num)]
        self.v_buffer = [torch.zeros(shape, dtype=self.store_dtype, device=device)
                         for _ in range(layer_num)]

    def get_kv_size_bytes(self):
        k = sum(t.numel() * t.element_size() for t in self.k_buffer)
        v = sum(t.numel() * t.element_size() for t in self.v_buffer)
        return k, v

    def _buf(self, bufs, layer_id: int):
        if self.layer_transfer_counter is not None:
            self.layer_transfer_counter.wait_until(layer_id - self.start_layer)
        t = bufs[layer_id - self.start_layer]
        return t.view(self.dtype) if self.store_dtype != self.dtype else t

    def get_key_buffer(self, layer_id: int):
        return self._buf(self.k_buffer, layer_id)

    def get_value_buffer(self, layer_id: int):
        return self._buf(self.v_buffer, layer_id)

    def get_kv_buffer(self, layer_id: int):
        return self.get_key_buffer(layer_id), self.get_value_buffer(layer_id)

    def set_kv_buffer(self, layer, loc: torch.Tensor, cache_k: torch.Tensor, cache_v: torch.Tensor,
                      k_scale: Optional[float] = None, v_scale: Optional[float] = None):
        idx = layer.layer_id - self.start_layer
        if cache_k.dtype != self.dtype:
            # FP8 pool: divide by the scale, cast and scatter in ONE kernel, bit-exact with the
            # reference's `div_` -> `.to(fp8)` -> index_put (memory_pool.py:390-412)
            if not _lib.is_fp8_kv(self.dtype):
                raise _lib.B200Error(f"KV cache dtype {self.dtype} is not supported "
                                     "(bf16/f16, float8_e4m3fn or float8_e5m2)")
            if (k_scale is None) != (v_scale is None):
                raise _lib.B200Error("set_kv_buffer: k_scale and v_scale must be given together")
            kv_write_fp8(cache_k, cache_v, self.k_buffer[idx], self.v_buffer[idx], loc,
                         k_scale, v_scale, self.dtype)
            return
        kv_write(cache_k, cache_v, self.k_buffer[idx], self.v_buffer[idx], loc)


def _scale_value(s) -> float:
    return float(s) if s is not None else 1.0


def kv_write_fp8(k: torch.Tensor, v: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor,
                 loc: torch.Tensor, k_scale=None, v_scale=None,
                 pool_dtype=torch.float8_e4m3fn) -> None:
    """Quantising scatter into an e4m3 (or e5m2) pool stored as bytes (b200_kv_write_fp8).  Scales
    are host floats (RadixAttention.k_scale_float / v_scale_float) so no device sync is needed."""
    if not k.is_cuda:
        raise _lib.B200Error("kv_write_fp8: tensors must live on a CUDA device (no CPU path)")
    num_tokens = loc.numel()
    if num_tokens == 0:
        return
    k = k.reshape(num_tokens, -1) if k.dim() != 2 else k
    v = v.reshape(num_tokens, -1) if v.dim() != 2 else v
    if k.stride(-1) != 1:
        k = k.contiguous()
    if v.stride(-1) != 1:
        v = v.contiguous()
    if loc.dtype != torch.int64:
        loc = loc.to(torch.int64)
    row = k.shape[1]
    if v.shape[1] != row or k_cache[0].numel() != row or v_cache[0].numel() != row \
            or k_cache.element_size() != 1 or not k_cache.is_contiguous():
        raise _lib.B200Error("kv_write_fp8: row does not match the cache row layout")
    rc = _lib.load().b200_kv_write_fp8(
        k.data_ptr(), v.data_ptr(), k.stride(0), v.stride(0), k_cache.data_ptr(),
        v_cache.data_ptr(), loc.data_ptr(), num_tokens, row, _scale_value(k_scale),
        _scale_value(v_scale), 0 if k_scale is None else 1, _lib.kv_dtype_code(k.dtype, pool_dtype),
        _lib.current_stream())
    _lib.check(rc, "b200_kv_write_fp8")


def kv_write(k: torch.Tensor, v: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor,
             loc: torch.Tensor) -> None:
    """k_cache[loc] = k ; v_cache[loc] = v  as one bit-exact row-copy kernel (b200_kv_write)."""
    if not k.is_cuda:
        raise _lib.B200Error("kv_write: tensors must live on a CUDA device (no CPU path)")
    num_tokens = loc.numel()
    if num_tokens == 0:
        return
    k = k.reshape(num_tokens, -1) if k.dim() != 2 else k
    v = v.reshape(num_tokens, -1) if v.dim() != 2 else v
    if k.stride(-1) != 1:
        k = k.contiguous()
    if v.stride(-1) != 1:
        v = v.contiguous()
    if loc.dtype != torch.int64:
        loc = loc.to(torch.int64)
    es_k, es_v = k.element_size(), v.element_size()
    k_row, v_row = k.shape[1] * es_k, v.shape[1] * es_v
    if k_cache[0].numel() * k_cache.element_size() != k_row or not k_cache.is_contiguous():
        raise _lib.B200Error("kv_write: k row does not match the cache row layout")
    if v_cache[0].numel() * v_cache.element_size() != v_row or not v_cache.is_contiguous():
        raise _lib.B200Error("kv_write: v row does not match the cache row layout")
    lib = _lib.load()
    rc = lib.b200_kv_write(k.data_ptr(), v.data_ptr(), k.stride(0) * es_k, v.stride(0) * es_v,
                           k_cache.data_ptr(), v_cache.data_ptr(), loc.data_ptr(), num_tokens,
                           k_row, v_row, _lib.current_stream())
    _lib.check(rc, "b200_kv_write")
