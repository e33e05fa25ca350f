"""GPU parity of the decode path (C-ABI -> CUDA) against the oracle and the reference goldens."""
import os

import pytest
import torch

from oracle import scenarios

import harness

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")

# bf16/f16 attention tolerance (SURVEY §8c note iii): max-abs <= 3e-2, relative L2 <= 1e-2
MAX_ABS, REL_L2 = 3e-2, 1e-2


def _check(case, num_kv_splits=0, logit_cap=0.0, golden=None, **kw):
    o_ref, k_ref, v_ref = harness.oracle_run(case, logit_cap=logit_cap)
    o, kb, vb = harness.b200_run(case, logit_cap=logit_cap, num_kv_splits=num_kv_splits, **kw)
    # cache-location writes are bit-exact, everywhere in the pool (nothing else may be touched)
    assert torch.equal(kb.view(torch.int16), k_ref.view(torch.int16))
    assert torch.equal(vb.view(torch.int16), v_ref.view(torch.int16))
    st = harness.error_stats(o, o_ref)
    assert st["max_abs"] <= MAX_ABS and st["rel_l2"] <= REL_L2, st
    if golden is not None:
        st = harness.error_stats(o, golden["o"])
        assert st["max_abs"] <= MAX_ABS and st["rel_l2"] <= REL_L2, ("vs reference golden", st)
    return st


@pytest.mark.parametrize("name", sorted(scenarios.DECODE_CASES))
@pytest.mark.parametrize("splits", [0, 1, 3])
def test_decode_matches_oracle_and_golden(name, splits):
    case = scenarios.make_attention_case(**scenarios.DECODE_CASES[name])
    g = torch.load(os.path.join(GOLDEN, f"attn_{name}.pt"), weights_only=False)
    _check(case, num_kv_splits=splits, golden=g)


@pytest.mark.parametrize("hq,hkv,d", [(32, 8, 128), (14, 2, 64), (8, 1, 128), (40, 8, 128),
                                      (6, 2, 128), (16, 1, 64), (3, 3, 64), (32, 2, 128),
                                      (24, 2, 128), (40, 2, 64), (64, 8, 128)])
def test_decode_geometries(hq, hkv, d):
    case = scenarios.make_attention_case(seed=hq * 100 + d, hq=hq, hkv=hkv, d=d,
                                         seq_lens=[1, 2, 15, 16, 17, 63, 64, 65, 300, 1025])
    for splits in (0, 2, 7):
        _check(case, num_kv_splits=splits)


def test_decode_long_context_many_splits():
    case = scenarios.make_attention_case(seed=5, hq=32, hkv=8, d=128, seq_lens=[4097, 2500, 8191, 1])
    for splits in (1, 8, 16, 33):
        _check(case, num_kv_splits=splits)


def test_decode_vectorised_slot_fetch_paths():
    """req_to_token row stride % 4 == 0 -> 16-byte slot-id loads for full tiles, scalar loads on
    the ragged tail; odd strides -> scalar everywhere.  Both must give identical results."""
    for lens in ([4096, 2500, 1, 48, 96, 97, 49, 47], [4093, 16, 32, 33]):
        case = scenarios.make_attention_case(seed=9, hq=32, hkv=8, d=128, seq_lens=lens)
        assert (case.req_to_token.shape[1] % 4 == 0) == (lens[0] == 4096)
        for splits in (0, 1, 5):
            _check(case, num_kv_splits=splits)


def test_decode_distribution_with_offset():
    # N(0.1, 0.2) inputs as in test/srt/test_triton_attention_kernels.py:67-90
    case = scenarios.make_attention_case(seed=6, hq=32, hkv=8, d=128, seq_lens=[777, 1300],
                                         mean=0.1, std=0.2)
    _check(case, num_kv_splits=4)


def test_decode_contiguous_qkv_and_f16():
    case = scenarios.make_attention_case(seed=7, hq=8, hkv=2, d=128, seq_lens=[100, 260],
                                         dtype=torch.float16)
    _check(case, num_kv_splits=2, strided_qkv=False)


def test_decode_logit_cap():
    case = scenarios.make_attention_case(seed=8, hq=8, hkv=2, d=128, seq_lens=[90, 200])
    _check(case, num_kv_splits=2, logit_cap=30.0)


def test_decode_int32_index_tensors():
    """req_pool_indices / seq_lens as int32 (what the reference's CudaGraphRunner allocates,
    cuda_graph_runner.py:250-253): the converted copies must stay alive until the launch — both are
    bs*8 bytes, so a freed temporary of the first would be overwritten by the second."""
    for lens in ([1, 63, 300, 1025, 700], [4097, 2500, 77]):
        case = scenarios.make_attention_case(seed=41, hq=32, hkv=8, d=128, seq_lens=lens)
        # request slots deliberately != lengths so a swapped pointer cannot pass by accident
        for splits in (0, 3):
            _check(case, num_kv_splits=splits, index_dtype=torch.int32)


@pytest.mark.parametrize("index_dtype", [torch.int64, torch.int32])
def test_decode_cuda_graph_capture_and_replay(index_dtype):
    """Decode step captured once, replayed after req_to_token / seq_lens / q change in place."""
    from sglang_security_b200.layers.attention.b200_backend import B200AttnBackend
    from sglang_security_b200.layers.radix_attention import RadixAttention
    from sglang_security_b200.model_executor.forward_batch_info import ForwardBatch, ForwardMode

    case_a = scenarios.make_attention_case(seed=31, hq=8, hkv=2, d=128, seq_lens=[50, 300, 129],
                                           pool_size=2000, max_reqs=8)
    case_b = scenarios.make_attention_case(seed=32, hq=8, hkv=2, d=128, seq_lens=[700, 2, 64],
                                           pool_size=2000, max_reqs=8)
    dev = "cuda"
    # one pool sized for the larger table
    ctx = max(case_a.req_to_token.shape[1], case_b.req_to_token.shape[1])
    for c in (case_a, case_b):
        pad = torch.zeros((c.req_to_token.shape[0], ctx), dtype=torch.int32)
        pad[:, :c.req_to_token.shape[1]] = c.req_to_token
        c.req_to_token = pad
    mr = harness.make_model_runner(case_a, dev)
    backend = B200AttnBackend(mr)
    layer = RadixAttention(case_a.hq, case_a.d, float(case_a.d) ** -0.5, case_a.hkv, 0)
    bs, t = case_a.bs, case_a.bs
    q = case_a.q.to(dev)
    k = case_a.k_new.to(dev)
    v = case_a.v_new.to(dev)
    req_pool_indices = case_a.req_pool_indices.to(dev).to(index_dtype)
    seq_lens = case_a.seq_lens.to(dev).to(index_dtype)
    out_cache_loc = case_a.out_cache_loc.to(dev)
    fb = ForwardBatch(forward_mode=ForwardMode.DECODE, batch_size=bs,
                      req_pool_indices=req_pool_indices, seq_lens=seq_lens,
                      out_cache_loc=out_cache_loc, seq_lens_sum=int(case_a.seq_lens.sum()),
                      req_to_token_pool=mr.req_to_token_pool,
                      token_to_kv_pool=mr.token_to_kv_pool, attn_backend=backend)
    backend.init_cuda_graph_state(max_bs=bs)
    backend.init_forward_metadata_capture_cuda_graph(bs, t, req_pool_indices, seq_lens, None,
                                                     ForwardMode.DECODE, None)
    # warm-up outside capture, then capture
    o = layer(q, k, v, fb)
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        o = layer(q, k, v, fb)
    graph.replay()
    torch.cuda.synchronize()
    ref_a, _, _ = harness.oracle_run(case_a)
    st = harness.error_stats(o.cpu(), ref_a)
    assert st["max_abs"] <= MAX_ABS and st["rel_l2"] <= REL_L2, st

    # new step: overwrite the persistent inputs in place, replay
    mr.req_to_token_pool.req_to_token.copy_(case_b.req_to_token)
    mr.token_to_kv_pool.k_buffer[0].copy_(case_b.k_cache)
    mr.token_to_kv_pool.v_buffer[0].copy_(case_b.v_cache)
    q.copy_(case_b.q); k.copy_(case_b.k_new); v.copy_(case_b.v_new)
    req_pool_indices.copy_(case_b.req_pool_indices)
    seq_lens.copy_(case_b.seq_lens)
    out_cache_loc.copy_(case_b.out_cache_loc)
    backend.init_forward_metadata_replay_cuda_graph(bs, req_pool_indices, seq_lens,
                                                    int(case_b.seq_lens.sum()), None,
                                                    ForwardMode.DECODE, None, case_b.seq_lens)
    graph.replay()
    torch.cuda.synchronize()
    ref_b, kb, vb = harness.oracle_run(case_b)
    st = harness.error_stats(o.cpu(), ref_b)
    assert st["max_abs"] <= MAX_ABS and st["rel_l2"] <= REL_L2, st
    assert torch.equal(mr.token_to_kv_pool.k_buffer[0].cpu().view(torch.int16), kb.view(torch.int16))


def test_decode_full_size_properties():
    """BASELINE config 2 geometry at full size (bs 64 x 4k, Hq 32 / Hkv 8 / D 128): checked through
    size-independent properties — (1) a request whose cached V rows are all the same vector must
    return exactly that vector (softmax weights sum to 1); (2) permuting the slot assignment of a
    request (same K/V content at different slots) leaves the output unchanged; (3) a sampled
    subset of requests agrees with the oracle."""
    dev = "cuda"
    torch.manual_seed(0)
    bs, ctx, hq, hkv, d = 64, 4096, 32, 8, 128
    size = bs * ctx
    from sglang_security_b200.mem_cache.memory_pool import MHATokenToKVPool, ReqToTokenPool
    from sglang_security_b200.layers.attention.b200_backend import B200AttnBackend
    from sglang_security_b200.layers.radix_attention import RadixAttention
    from sglang_security_b200.model_executor.forward_batch_info import ForwardBatch, ForwardMode
    from types import SimpleNamespace

    rtp = ReqToTokenPool(bs + 1, ctx + 4, dev)
    kvp = MHATokenToKVPool(size, 1, torch.bfloat16, hkv, d, 1, dev)
    kvp.k_buffer[0].normal_()
    kvp.v_buffer[0].normal_()
    perm = torch.randperm(size, device=dev) + 1
    rtp.req_to_token[:bs, :ctx] = perm.view(bs, ctx).to(torch.int32)
    # request 0: constant V rows
    const_v = torch.randn(hkv, d, device=dev, dtype=torch.bfloat16)
    kvp.v_buffer[0][rtp.req_to_token[0, :ctx].long()] = const_v
    # request 1 and 2: same content, different slots
    s1, s2 = rtp.req_to_token[1, :ctx].long(), rtp.req_to_token[2, :ctx].long()
    kvp.k_buffer[0][s2] = kvp.k_buffer[0][s1]
    kvp.v_buffer[0][s2] = kvp.v_buffer[0][s1]
    cfg = SimpleNamespace(num_attention_heads=hq, context_len=ctx, get_num_kv_heads=lambda tp: hkv)
    mr = SimpleNamespace(device=dev, req_to_token_pool=rtp, token_to_kv_pool=kvp, model_config=cfg,
                         tp_size=1, server_args=SimpleNamespace(b200_attention_num_kv_splits=0))
    backend = B200AttnBackend(mr)
    layer = RadixAttention(hq, d, d ** -0.5, hkv, 0)
    q = torch.randn(bs, hq * d, device=dev, dtype=torch.bfloat16)
    q[2] = q[1]
    seq_lens = torch.full((bs,), ctx, dtype=torch.int64, device=dev)
    fb = ForwardBatch(forward_mode=ForwardMode.DECODE, batch_size=bs,
                      req_pool_indices=torch.arange(bs, device=dev), seq_lens=seq_lens,
                      out_cache_loc=rtp.req_to_token[:bs, ctx - 1].long(), seq_lens_sum=bs * ctx,
                      req_to_token_pool=rtp, token_to_kv_pool=kvp, attn_backend=backend)
    backend.init_forward_metadata(fb)
    o = layer(q, None, None, fb, save_kv_cache=False).view(bs, hq, d)
    torch.cuda.synchronize()
    expect = const_v.repeat_interleave(hq // hkv, dim=0).float()
    assert (o[0].float() - expect).abs().max().item() <= 1e-2      # property 1 (bf16 rounding)
    assert torch.equal(o[1], o[2])                                 # property 2 (bit-identical)
    from oracle.attention import decode_attention_oracle
    pick = [5, 63]
    ref = decode_attention_oracle(
        q[pick].cpu().view(len(pick), hq, d), kvp.k_buffer[0].cpu(), kvp.v_buffer[0].cpu(),
        rtp.req_to_token.cpu(), torch.tensor(pick), seq_lens[pick].cpu(), d ** -0.5)
    st = harness.error_stats(o[pick].cpu(), ref)
    assert st["max_abs"] <= MAX_ABS and st["rel_l2"] <= REL_L2, st


# ---- fused tail: merge (+ o_proj input quantisation) by the last CTA of each request ---------------
@pytest.mark.parametrize("hq,hkv,d", [(32, 8, 128), (14, 2, 64), (8, 1, 128), (64, 8, 128)])
@pytest.mark.parametrize("splits", [1, 2, 5])
@pytest.mark.parametrize("in_kernel", [True, False])
def test_decode_fused_tail_equals_split_plus_merge_plus_quant(hq, hkv, d, splits, in_kernel):
    """The fused tail must reproduce the three-launch chain bit for bit: merged 16-bit row ==
    b200_attn_merge's, e4m3 bytes and fp32 scales == b200_fp8_quant_per_token of that row; three
    back-to-back calls prove the ticket counters re-arm themselves (ragged lengths incl. requests
    with fewer non-empty splits than the grid has)."""
    from sglang_security_b200.layers.quantization.fp8_kernel import sglang_per_token_quant_fp8

    case = scenarios.make_attention_case(seed=11 * hq + d + splits, hq=hq, hkv=hkv, d=d,
                                         seq_lens=[1, 47, 48, 49, 300, 1025, 700, 96], shared_prefix=16)
    o_plain, _, _ = harness.b200_run(case, num_kv_splits=splits, fused_tail=False)
    o_fused, _, _ = harness.b200_run(case, num_kv_splits=splits, fused_tail=True, repeat=3,
                                     tail_in_kernel=in_kernel)
    assert torch.equal(o_plain.view(torch.int16), o_fused.view(torch.int16))
    (qa, sa), _, _ = harness.b200_run(case, num_kv_splits=splits, fused_tail=True, fp8_out=True, repeat=3,
                                      tail_in_kernel=in_kernel)
    q_ref, s_ref = sglang_per_token_quant_fp8(o_plain.cuda())
    assert torch.equal(sa, s_ref.cpu())
    assert torch.equal(qa.view(torch.uint8), q_ref.cpu().view(torch.uint8))
    # and the unfused fallback of fp8_out (tail disabled) gives the same pair
    (qb, sb), _, _ = harness.b200_run(case, num_kv_splits=splits, fused_tail=False, fp8_out=True)
    assert torch.equal(sb, s_ref.cpu()) and torch.equal(qb.view(torch.uint8), q_ref.cpu().view(torch.uint8))


def test_decode_fused_tail_full_batch_against_oracle():
    """cfg-2 head geometry, 64 requests, heuristic split count: fused tail vs the oracle."""
    case = scenarios.make_attention_case(seed=77, hq=32, hkv=8, d=128,
                                         seq_lens=[257 + 13 * i for i in range(64)], shared_prefix=128)
    o_ref, _, _ = harness.oracle_run(case)
    o, _, _ = harness.b200_run(case, num_kv_splits=4, fused_tail=True, repeat=2, tail_in_kernel=True)
    st = harness.error_stats(o, o_ref)
    assert st["max_abs"] <= MAX_ABS and st["rel_l2"] <= REL_L2, st


def test_decode_fused_tail_ticket_sizing():
    from sglang_security_b200 import _lib

    lib = _lib.load()
    assert lib.b200_attn_decode_fused_ticket_bytes(64, 32, 8) == 64 * (1 + 8) * 4
    assert lib.b200_attn_decode_fused_ticket_bytes(3, 40, 2) == 3 * (1 + 2 * 2) * 4   # group 20: two head tiles
    assert lib.b200_attn_decode_fused_ticket_bytes(1, 7, 2) == 0
