"""CPU: the scheduler-index oracle against the fixtures produced by the reference's own torch
implementations (oracle/gen_golden_sched_index.py), plus allocator invariants."""
import os

import torch

from oracle import sched_index as O

GOLD = os.path.join(os.path.dirname(__file__), "golden", "sched_index.pt")


def test_oracle_matches_reference_fixtures():
    for c in torch.load(GOLD):
        last = O.get_last_loc_oracle(c["req_to_token"], c["req_pool_indices"], c["prefix_lens"])
        assert torch.equal(last, c["last_loc"])
        pos, start = O.compute_position_oracle(c["prefix_lens"], c["extend_seq_lens"])
        assert torch.equal(pos, c["positions"])
        assert torch.equal(start, c["extend_start_loc"].to(torch.int32))


def test_alloc_oracle_invariants():
    page = 16
    pre = torch.tensor([0, 5, 16, 31, 40])
    seq = torch.tensor([3, 40, 17, 32, 100])
    # last_loc of a prefix ending mid-page: any slot with (loc + 1) % page == pre % page
    last = torch.tensor([-1, 16 * 50 + 4, 16 * 51 + 15, 16 * 52 + 14, 16 * 53 + 7])
    free = torch.arange(100, 160)
    out, used = O.alloc_extend_oracle(pre, seq, last, free, page)
    assert out.numel() == int((seq - pre).sum()) and out.unique().numel() == out.numel()
    assert used == sum((int(s) + page - 1) // page - (int(p) + page - 1) // page
                       for p, s in zip(pre, seq))
    # request 1 continues its page, then takes whole pages
    seg = out[3:3 + 35]
    assert seg[0] == last[1] + 1 and (seg[11] % page) == 0
    dec, used_d = O.alloc_decode_oracle(torch.tensor([17, 18, 33]), torch.tensor([15, 16 * 9, 16 * 7 + 15]),
                                        torch.tensor([70, 71]), page)
    assert dec.tolist() == [70 * 16, 16 * 9 + 1, 71 * 16] and used_d == 2


ALLOC_GOLD = os.path.join(os.path.dirname(__file__), "golden", "sched_alloc.pt")


def test_alloc_oracle_matches_reference_triton_kernels():
    """tests/golden/sched_alloc.pt = outputs of the reference's OWN alloc_extend_kernel /
    alloc_decode_kernel (paged_allocator.py:29-155), run unmodified through Triton's CPU interpreter
    (oracle/gen_golden_alloc.py): page sizes 1 / 16 / 32 / 64, empty and page-aligned prefixes,
    extends ending on a page boundary, one-token extends."""
    cases = torch.load(ALLOC_GOLD)
    assert len(cases) >= 6
    for c in cases:
        page = c["page_size"]
        out, used = O.alloc_extend_oracle(c["prefix_lens"], c["seq_lens"], c["last_loc"], c["free_pages"], page)
        assert torch.equal(out, c["extend_out_indices"]), page
        assert used == c["extend_num_new_pages"]
        assert int((c["seq_lens"] - c["prefix_lens"]).sum()) == c["extend_sum_extend_lens"]
        dec, used_d = O.alloc_decode_oracle(c["decode_seq_lens"], c["decode_last_loc"],
                                            c["decode_free_pages"], page)
        assert torch.equal(dec, c["decode_out_indices"]), page
        assert used_d == c["decode_num_new_pages"]


def test_index_oracles_match_reference_triton_kernels():
    """tests/golden/triton_index.pt = outputs of the reference's own create_flashinfer_kv_indices_triton
    (layers/attention/utils.py:5-41) and write_req_to_token_pool_triton (schedule_batch.py:1789-1823) run
    through Triton's CPU interpreter (oracle/gen_golden_triton_index.py)."""
    from oracle.pools import kv_indices_oracle

    gold = torch.load(os.path.join(os.path.dirname(__file__), "golden", "triton_index.pt"))
    for c in gold["kv_indices"]:
        indptr, flat = kv_indices_oracle(c["req_to_token"], c["req_pool_indices"], c["lens"], c["kv_start"])
        assert torch.equal(indptr, c["kv_indptr"]) and torch.equal(flat, c["kv_indices"])
    for c in gold["write_req_to_token"]:
        got = O.write_req_to_token_oracle(c["req_to_token_before"].clone(), c["req_pool_indices"], c["pre_lens"],
                                          c["seq_lens"], c["extend_lens"], c["out_cache_loc"])
        assert torch.equal(got, c["req_to_token_after"])
