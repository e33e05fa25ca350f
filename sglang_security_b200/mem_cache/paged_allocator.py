"""Page-aligned KV slot allocator and the scheduler-side index helpers, on the B200 index kernels.

Host-side mirror of the reference (same names, arguments and return conventions):
  PagedTokenToKVPoolAllocator  python/sglang/srt/mem_cache/paged_allocator.py:157-320
  write_req_to_token_pool      managers/schedule_batch.py:1272-1291, 1789-1823
  get_last_loc                 managers/schedule_batch.py:1826-1893
  compute_position             model_executor/forward_batch_info.py:639-692
The pool stays TOKEN-indexed (`page * page_size + offset`, paged_allocator.py:210-215), so the
attention kernels need no page-size awareness: pages are aligned runs of slot ids.
The allocation policy (free list order, what is popped when) is the reference's; only the index
arithmetic runs in our kernels (csrc/sched_index.cu).  There is no CPU path.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from .. import _lib


def _i64(t: torch.Tensor) -> torch.Tensor:
    return t if t.dtype == torch.int64 else t.to(torch.int64)


def _need_cuda(t: torch.Tensor, what: str) -> None:
    if not t.is_cuda:
        raise _lib.B200Error(f"{what}: tensors must live on a CUDA device (no CPU path)")


def write_req_to_token_pool(req_to_token: torch.Tensor, req_pool_indices: torch.Tensor,
                            pre_lens: torch.Tensor, seq_lens: torch.Tensor,
                            extend_lens: torch.Tensor, out_cache_loc: torch.Tensor) -> None:
    """req_to_token[req, pre:seq] = the request's slice of out_cache_loc (in place)."""
    _need_cuda(req_to_token, "write_req_to_token_pool")
    if req_to_token.dtype != torch.int32 or req_to_token.stride(1) != 1:
        raise _lib.B200Error("write_req_to_token_pool: req_to_token must be int32, row-major")
    bs = req_pool_indices.numel()
    # converted copies are bound to locals: a temporary would be freed (and its block handed to the
    # next conversion) before the launch reads it
    rpi, pre, seq, ext, loc = (_i64(req_pool_indices), _i64(pre_lens), _i64(seq_lens),
                               _i64(extend_lens), _i64(out_cache_loc))
    rc = _lib.load().b200_write_req_to_token(
        req_to_token.data_ptr(), req_to_token.stride(0), rpi.data_ptr(), pre.data_ptr(),
        seq.data_ptr(), ext.data_ptr(), loc.data_ptr(), bs, _lib.current_stream())
    _lib.check(rc, "b200_write_req_to_token")


def get_last_loc(req_to_token: torch.Tensor, req_pool_indices_tensor: torch.Tensor,
                 prefix_lens_tensor: torch.Tensor) -> torch.Tensor:
    _need_cuda(req_to_token, "get_last_loc")
    n = prefix_lens_tensor.numel()
    result = torch.empty(n, dtype=torch.int64, device=req_to_token.device)
    rpi, pre = _i64(req_pool_indices_tensor), _i64(prefix_lens_tensor)
    rc = _lib.load().b200_get_last_loc(
        req_to_token.data_ptr(), req_to_token.stride(0), rpi.data_ptr(), pre.data_ptr(),
        result.data_ptr(), n, _lib.current_stream())
    _lib.check(rc, "b200_get_last_loc")
    return result.to(prefix_lens_tensor.dtype)


def compute_position(extend_prefix_lens: Optional[torch.Tensor], extend_seq_lens: torch.Tensor,
                     extend_seq_lens_sum: int) -> Tuple[torch.Tensor, torch.Tensor]:
    """-> (positions int64 [sum], extend_start_loc int32 [bs])."""
    _need_cuda(extend_seq_lens, "compute_position")
    bs = extend_seq_lens.numel()
    dev = extend_seq_lens.device
    positions = torch.empty(extend_seq_lens_sum, dtype=torch.int64, device=dev)
    start = torch.empty(bs, dtype=torch.int32, device=dev)
    has_prefix = extend_prefix_lens is not None and extend_prefix_lens.numel() == bs
    pre = extend_prefix_lens.to(torch.int32) if has_prefix else None
    ext = extend_seq_lens.to(torch.int32)
    rc = _lib.load().b200_compute_position(
        positions.data_ptr(), start.data_ptr(), _lib.ptr(pre), ext.data_ptr(), bs,
        _lib.current_stream())
    _lib.check(rc, "b200_compute_position")
    return positions, start


class PagedTokenToKVPoolAllocator:
    """Allocator whose per-request output is page-aligned (paged_allocator.py:157-320)."""

    def __init__(self, size: int, page_size: int, dtype: torch.dtype, device, kvcache=None):
        self.size = size
        self.dtype = dtype
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise _lib.B200Error("PagedTokenToKVPoolAllocator needs a CUDA device (no CPU path)")
        self.page_size = page_size
        self.num_pages = size // page_size
        self._kvcache = kvcache
        self.ret_values = torch.empty((), dtype=torch.int64, device=self.device)
        self.clear()

    def available_size(self) -> int:
        return len(self.free_pages) * self.page_size

    def get_kvcache(self):
        return self._kvcache

    def alloc(self, need_size: int):
        num_pages = need_size // self.page_size
        if num_pages > len(self.free_pages):
            return None
        out_pages = self.free_pages[:num_pages]
        self.free_pages = self.free_pages[num_pages:]
        return (out_pages[:, None] * self.page_size
                + torch.arange(self.page_size, device=self.device)).reshape(-1)

    def alloc_extend(self, prefix_lens: torch.Tensor, seq_lens: torch.Tensor,
                     last_loc: torch.Tensor, extend_num_tokens: int):
        bs = len(prefix_lens)
        out_indices = torch.empty((extend_num_tokens,), dtype=torch.int64, device=self.device)
        pre, seq, last = _i64(prefix_lens), _i64(seq_lens), _i64(last_loc)
        rc = _lib.load().b200_alloc_extend(
            pre.data_ptr(), seq.data_ptr(), last.data_ptr(),
            self.free_pages.data_ptr(), out_indices.data_ptr(), self.ret_values.data_ptr(), bs,
            self.page_size, _lib.current_stream())
        _lib.check(rc, "b200_alloc_extend")
        merged_value = self.ret_values.item()
        num_new_pages = merged_value >> 32
        if num_new_pages > len(self.free_pages):
            return None
        self.free_pages = self.free_pages[num_new_pages:]
        return out_indices

    def alloc_decode(self, seq_lens: torch.Tensor, last_loc: torch.Tensor):
        bs = len(seq_lens)
        out_indices = torch.empty((bs,), dtype=torch.int64, device=self.device)
        seq, last = _i64(seq_lens), _i64(last_loc)
        rc = _lib.load().b200_alloc_decode(
            seq.data_ptr(), last.data_ptr(), self.free_pages.data_ptr(),
            out_indices.data_ptr(), self.ret_values.data_ptr(), bs, self.page_size,
            _lib.current_stream())
        _lib.check(rc, "b200_alloc_decode")
        num_new_pages = self.ret_values.item()
        if num_new_pages > len(self.free_pages):
            return None
        self.free_pages = self.free_pages[num_new_pages:]
        return out_indices

    def free(self, free_index: torch.Tensor):
        if free_index.numel() == 0:
            return
        if self.is_not_in_free_group:
            free_page_indices = torch.unique(free_index // self.page_size)
            self.free_pages = torch.cat((free_page_indices, self.free_pages))
        else:
            self.free_group.append(free_index)

    def free_group_begin(self):
        self.is_not_in_free_group = False
        self.free_group = []

    def free_group_end(self):
        self.is_not_in_free_group = True
        if self.free_group:
            self.free(torch.cat(self.free_group))

    def backup_state(self):
        return self.free_pages

    def restore_state(self, free_pages):
        self.free_pages = free_pages

    def clear(self):
        # page 0 holds the padded slot 0 used for dummy writes of padded tokens
        self.free_pages = torch.arange(1, self.num_pages + 1, dtype=torch.int64,
                                       device=self.device)
        self.is_not_in_free_group = True
        self.free_group = []
