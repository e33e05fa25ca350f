"""TEST INFRASTRUCTURE ONLY (CPU oracle; never imported by the product path).

Python-loop restatements of the reference's scheduler-side index kernels:
  write_req_to_token_oracle  the non-Triton branch of ScheduleBatch.prepare_for_extend
                             (managers/schedule_batch.py:1284-1291) == write_req_to_token_pool_triton
                             (:1789-1823)
  get_last_loc_oracle        get_last_loc_torch (managers/schedule_batch.py:1838-1847)
  compute_position_oracle    compute_position_torch (model_executor/forward_batch_info.py:694-709)
  alloc_extend_oracle /      alloc_extend_kernel / alloc_decode_kernel
  alloc_decode_oracle        (mem_cache/paged_allocator.py:29-155), one request at a time

Pinned by: tests/golden/sched_index.pt = outputs of the reference's own get_last_loc_torch and
compute_position_torch (source extracted unmodified from the reference files by
oracle/gen_golden_sched_index.py), and tests/golden/sched_alloc.pt = outputs of the reference's own
Triton allocator kernels (source extracted unmodified, executed on CPU tensors through Triton's
interpreter by oracle/gen_golden_alloc.py), plus invariants (unique slots, page alignment,
continuation of the last page).
"""
from __future__ import annotations

import torch


def write_req_to_token_oracle(req_to_token, req_pool_indices, pre_lens, seq_lens, extend_lens,
                              out_cache_loc):
    pt = 0
    for i in range(len(req_pool_indices)):
        n = int(seq_lens[i]) - int(pre_lens[i])
        req_to_token[int(req_pool_indices[i]), int(pre_lens[i]):int(seq_lens[i])] = \
            out_cache_loc[pt:pt + n].to(req_to_token.dtype)
        pt += int(extend_lens[i])
    return req_to_token


def get_last_loc_oracle(req_to_token, req_pool_indices, prefix_lens):
    out = torch.full_like(prefix_lens, -1)
    for i in range(len(prefix_lens)):
        if int(prefix_lens[i]) > 0:
            out[i] = req_to_token[int(req_pool_indices[i]), int(prefix_lens[i]) - 1]
    return out


def compute_position_oracle(extend_prefix_lens, extend_seq_lens):
    pos, start, acc = [], [], 0
    for b in range(len(extend_seq_lens)):
        pre = int(extend_prefix_lens[b]) if extend_prefix_lens is not None else 0
        n = int(extend_seq_lens[b])
        pos.extend(range(pre, pre + n))
        start.append(acc)
        acc += n
    return torch.tensor(pos, dtype=torch.int64), torch.tensor(start, dtype=torch.int32)


def _pages(n, page):
    return (n + page - 1) // page


def alloc_extend_oracle(pre_lens, seq_lens, last_loc, free_pages, page):
    """-> (out_indices int64 [sum extend], num_new_pages)."""
    out, used = [], 0
    for b in range(len(seq_lens)):
        pre, seq, ll = int(pre_lens[b]), int(seq_lens[b]), int(last_loc[b])
        new_pages = _pages(seq, page) - _pages(pre, page)
        pre_up = _pages(pre, page) * page
        n1 = min(seq, pre_up) - pre
        out.extend(ll + 1 + i for i in range(n1))
        if pre + n1 < seq:
            n2 = seq // page * page - pre_up
            out.extend(int(free_pages[used + i // page]) * page + i % page for i in range(n2))
            if pre + n1 + n2 < seq:
                n3 = seq - seq // page * page
                lp = int(free_pages[used + new_pages - 1])
                out.extend(lp * page + i for i in range(n3))
        used += new_pages
    return torch.tensor(out, dtype=torch.int64), used


def alloc_decode_oracle(seq_lens, last_loc, free_pages, page):
    out, used = [], 0
    for b in range(len(seq_lens)):
        seq = int(seq_lens[b])
        new = _pages(seq, page) - _pages(seq - 1, page)
        out.append(int(last_loc[b]) + 1 if new == 0 else int(free_pages[used]) * page)
        used += new
    return torch.tensor(out, dtype=torch.int64), used
