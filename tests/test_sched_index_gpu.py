"""GPU parity (bit-exact) of the scheduler-side index kernels and of the paged allocator built on
them against oracle/sched_index.py and the reference-generated fixtures; then paged decode attention
on page-aligned slots (page_size 16 / 64), which the token-indexed kernels serve unchanged."""
import os

import pytest
import torch

from oracle import sched_index as O

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden", "sched_index.pt")


def test_get_last_loc_and_positions_match_reference_fixtures():
    from sglang_security_b200.mem_cache.paged_allocator import compute_position, get_last_loc

    for c in torch.load(GOLD):
        last = get_last_loc(c["req_to_token"].cuda(), c["req_pool_indices"].cuda(),
                            c["prefix_lens"].cuda())
        assert torch.equal(last.cpu(), c["last_loc"])
        pos, start = compute_position(c["prefix_lens"].cuda().to(torch.int32),
                                      c["extend_seq_lens"].cuda(), int(c["extend_seq_lens"].sum()))
        assert torch.equal(pos.cpu(), c["positions"])
        assert torch.equal(start.cpu(), c["extend_start_loc"].to(torch.int32))
        pos0, _ = compute_position(None, c["extend_seq_lens"].cuda(), int(c["extend_seq_lens"].sum()))
        ref0, _ = O.compute_position_oracle(None, c["extend_seq_lens"])
        assert torch.equal(pos0.cpu(), ref0)


@pytest.mark.parametrize("bs", [1, 9, 300, 1500])
def test_write_req_to_token_bit_exact(bs):
    from sglang_security_b200.mem_cache.paged_allocator import write_req_to_token_pool

    g = torch.Generator().manual_seed(bs)
    ctx = 700
    pre = torch.randint(0, 300, (bs,), generator=g)
    ext = torch.randint(1, 400, (bs,), generator=g)
    seq = pre + ext
    rpi = torch.randperm(bs + 5, generator=g)[:bs]
    loc = torch.randint(1, 1 << 30, (int(ext.sum()),), generator=g)
    table = torch.randint(0, 100, (bs + 5, ctx), generator=g, dtype=torch.int32)
    ref = O.write_req_to_token_oracle(table.clone(), rpi, pre, seq, ext, loc)
    dev = table.cuda()
    write_req_to_token_pool(dev, rpi.cuda(), pre.cuda(), seq.cuda(), ext.cuda(), loc.cuda())
    assert torch.equal(dev.cpu(), ref)


def test_index_helpers_accept_int32_inputs():
    """int32 index tensors (the dtype the reference's scheduler uses for seq_lens / prefix lens on
    some paths): every converted copy must outlive its launch — five same-sized conversions in a row
    used to land in the same allocator block."""
    from sglang_security_b200.mem_cache.paged_allocator import (PagedTokenToKVPoolAllocator,
                                                                get_last_loc,
                                                                write_req_to_token_pool)

    g = torch.Generator().manual_seed(3)
    bs, ctx = 48, 700
    pre = torch.randint(0, 300, (bs,), generator=g)
    ext = torch.randint(1, 400, (bs,), generator=g)
    seq = pre + ext
    rpi = torch.randperm(bs + 5, generator=g)[:bs]
    loc = torch.randint(1, 1 << 30, (int(ext.sum()),), generator=g)
    table = torch.randint(0, 100, (bs + 5, ctx), generator=g, dtype=torch.int32)
    ref = O.write_req_to_token_oracle(table.clone(), rpi, pre, seq, ext, loc)
    i32 = lambda t: t.to(torch.int32).cuda()
    dev = table.cuda()
    write_req_to_token_pool(dev, i32(rpi), i32(pre), i32(seq), i32(ext), loc.cuda())
    assert torch.equal(dev.cpu(), ref)
    last = get_last_loc(dev, i32(rpi), i32(pre))
    last_ref = get_last_loc(dev, rpi.cuda(), pre.cuda())
    assert torch.equal(last.cpu().long(), last_ref.cpu())
    for page in (16, 64):
        a32 = PagedTokenToKVPoolAllocator(64 * 1024, page, torch.bfloat16, "cuda")
        a64 = PagedTokenToKVPoolAllocator(64 * 1024, page, torch.bfloat16, "cuda")
        out32 = a32.alloc_extend(i32(pre), i32(seq), i32(last_ref), int(ext.sum()))
        out64 = a64.alloc_extend(pre.cuda(), seq.cuda(), last_ref, int(ext.sum()))
        assert torch.equal(out32, out64) and torch.equal(a32.free_pages, a64.free_pages)
        d32 = a32.alloc_decode(i32(seq + 1), i32(last_ref))
        d64 = a64.alloc_decode((seq + 1).cuda(), last_ref)
        assert torch.equal(d32, d64) and torch.equal(a32.free_pages, a64.free_pages)


@pytest.mark.parametrize("page", [1, 16, 64])
def test_paged_allocator_extend_then_decode(page):
    from sglang_security_b200.mem_cache.paged_allocator import (PagedTokenToKVPoolAllocator,
                                                                get_last_loc,
                                                                write_req_to_token_pool)

    g = torch.Generator().manual_seed(page)
    bs, ctx = 37, 600
    size = 64 * 1024
    alloc = PagedTokenToKVPoolAllocator(size, page, torch.bfloat16, "cuda")
    table = torch.zeros((bs, ctx), dtype=torch.int32, device="cuda")
    rpi = torch.arange(bs, device="cuda")
    # step 1: extend from scratch; step 2: extend with the first result as cached prefix
    pre = torch.zeros(bs, dtype=torch.int64)
    for step in range(2):
        ext = torch.randint(1, 200, (bs,), generator=g)
        seq = pre + ext
        last = get_last_loc(table, rpi, pre.cuda())
        free_before = alloc.free_pages.cpu()
        out = alloc.alloc_extend(pre.cuda(), seq.cuda(), last, int(ext.sum()))
        ref, used = O.alloc_extend_oracle(pre, seq, last.cpu(), free_before, page)
        assert torch.equal(out.cpu(), ref)
        assert torch.equal(alloc.free_pages.cpu(), free_before[used:])
        write_req_to_token_pool(table, rpi, pre.cuda(), seq.cuda(), ext.cuda(), out)
        pre = seq
    # every request's slots: unique, and position p of a request sits at offset p % page of a page
    t = table.cpu()
    for b in range(bs):
        row = t[b, :int(pre[b])].long()
        assert row.unique().numel() == row.numel()
        assert torch.equal(row % page, torch.arange(int(pre[b])) % page)
    # decode steps
    for _ in range(3):
        seq = pre + 1
        last = get_last_loc(table, rpi, pre.cuda())
        free_before = alloc.free_pages.cpu()
        out = alloc.alloc_decode(seq.cuda(), last)
        ref, used = O.alloc_decode_oracle(seq, last.cpu(), free_before, page)
        assert torch.equal(out.cpu(), ref)
        assert torch.equal(alloc.free_pages.cpu(), free_before[used:])
        write_req_to_token_pool(table, rpi, pre.cuda(), seq.cuda(), torch.ones(bs, dtype=torch.int64).cuda(), out)
        pre = seq
    # freeing everything returns every page exactly once
    all_slots = torch.cat([table[b, :int(pre[b])].long() for b in range(bs)])
    alloc.free(all_slots)
    assert alloc.free_pages.unique().numel() == alloc.num_pages == alloc.free_pages.numel()


@pytest.mark.parametrize("page", [16, 64])
def test_decode_attention_on_paged_slots(page):
    """page_size > 1: slots handed out by the paged allocator, attention through the backend."""
    import harness
    from oracle import scenarios
    from sglang_security_b200.mem_cache.paged_allocator import (PagedTokenToKVPoolAllocator,
                                                                get_last_loc)

    seq_lens = [1, 17, 64, 65, 300, 1000]
    case = scenarios.make_attention_case(seed=page, hq=32, hkv=8, d=128, seq_lens=seq_lens,
                                         pool_size=4096)
    alloc = PagedTokenToKVPoolAllocator(4096 - 2 * page, page, torch.bfloat16, "cuda")  # max slot < pool rows
    bs = len(seq_lens)
    pre = torch.zeros(bs, dtype=torch.int64)
    seq = torch.tensor(seq_lens)
    last = torch.full((bs,), -1, dtype=torch.int64)
    out = alloc.alloc_extend(pre.cuda(), seq.cuda(), last.cuda(), int(seq.sum())).cpu()
    case.req_to_token.zero_()
    pt = 0
    locs = []
    for b in range(bs):
        case.req_to_token[case.req_pool_indices[b], :seq_lens[b]] = out[pt:pt + seq_lens[b]].to(torch.int32)
        locs.append(out[pt + seq_lens[b] - 1])
        pt += seq_lens[b]
    case.out_cache_loc = torch.stack(locs)
    o_ref, k_ref, v_ref = harness.oracle_run(case)
    o, kb, vb = harness.b200_run(case)
    assert torch.equal(kb.view(torch.int16), k_ref.view(torch.int16))
    st = harness.error_stats(o, o_ref)
    assert st["max_abs"] <= 3e-2 and st["rel_l2"] <= 1e-2, st


def test_alloc_kernels_match_reference_triton_goldens():
    """b200_alloc_extend / b200_alloc_decode against tests/golden/sched_alloc.pt = the outputs of the
    reference's own Triton allocator kernels (oracle/gen_golden_alloc.py), bit-exact, including the
    pages consumed."""
    import os

    from sglang_security_b200.mem_cache.paged_allocator import PagedTokenToKVPoolAllocator

    cases = torch.load(os.path.join(os.path.dirname(__file__), "golden", "sched_alloc.pt"))
    for c in cases:
        page = c["page_size"]
        alloc = PagedTokenToKVPoolAllocator(8192 * page, page, torch.bfloat16, "cuda")
        alloc.free_pages = c["free_pages"].cuda()
        out = alloc.alloc_extend(c["prefix_lens"].cuda(), c["seq_lens"].cuda(), c["last_loc"].cuda(),
                                 int(c["extend_sum_extend_lens"]))
        assert torch.equal(out.cpu(), c["extend_out_indices"]), page
        assert torch.equal(alloc.free_pages.cpu(), c["free_pages"][c["extend_num_new_pages"]:])
        assert torch.equal(alloc.free_pages.cpu(), c["decode_free_pages"])
        dec = alloc.alloc_decode(c["decode_seq_lens"].cuda(), c["decode_last_loc"].cuda())
        assert torch.equal(dec.cpu(), c["decode_out_indices"]), page
        assert alloc.free_pages.numel() == c["decode_free_pages"].numel() - c["decode_num_new_pages"]


def test_index_kernels_match_reference_triton_goldens():
    """b200_kv_indices / b200_write_req_to_token against the outputs of the reference's own Triton kernels
    (tests/golden/triton_index.pt, oracle/gen_golden_triton_index.py): bit-exact."""
    import os

    from sglang_security_b200 import _lib
    from sglang_security_b200.mem_cache.paged_allocator import write_req_to_token_pool

    gold = torch.load(os.path.join(os.path.dirname(__file__), "golden", "triton_index.pt"))
    lib = _lib.load()
    for c in gold["kv_indices"]:
        rtt = c["req_to_token"].cuda()
        rpi = c["req_pool_indices"].to(torch.int64).cuda()
        lens = c["lens"].to(torch.int64).cuda()
        indptr = c["kv_indptr"].cuda()
        start = None if c["kv_start"] is None else c["kv_start"].cuda()
        out = torch.full((int(c["kv_indptr"][-1]),), -1, dtype=torch.int32, device="cuda")
        rc = lib.b200_kv_indices(rtt.data_ptr(), rtt.stride(0), rpi.data_ptr(), lens.data_ptr(),
                                 indptr.data_ptr(), _lib.ptr(start), out.data_ptr(), rpi.numel(),
                                 _lib.current_stream())
        _lib.check(rc, "b200_kv_indices")
        assert torch.equal(out.cpu(), c["kv_indices"])
    for c in gold["write_req_to_token"]:
        table = c["req_to_token_before"].cuda()
        write_req_to_token_pool(table, c["req_pool_indices"].cuda(), c["pre_lens"].cuda(), c["seq_lens"].cuda(),
                                c["extend_lens"].cuda(), c["out_cache_loc"].cuda())
        assert torch.equal(table.cpu(), c["req_to_token_after"])
