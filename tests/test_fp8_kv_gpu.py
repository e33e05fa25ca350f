"""GPU parity of the FP8 (e4m3) KV-cache path — SURVEY §8(f) rank 1: quantising set_kv_buffer
(bit-exact with the reference's div_ -> .to(float8_e4m3fn) -> index_put, memory_pool.py:390-412,
restated in oracle/pools.py) and paged decode attention over the e4m3 pool
(oracle/attention.py: unmodified torch_native math on the de-quantised cache with the FlashInfer
backend's k_scale / v_scale semantics).  Tolerances: max-abs <= 3e-2, rel-l2 <= 1e-2."""
from types import SimpleNamespace

import pytest
import torch

from oracle import attention as OA
from oracle import scenarios
from oracle.pools import OracleMHATokenToKVPool

pytestmark = pytest.mark.gpu
FP8 = torch.float8_e4m3fn


def _layer(hq, hkv, d, k_scale=None, v_scale=None, logit_cap=0.0):
    from sglang_security_b200.layers.radix_attention import RadixAttention

    layer = RadixAttention(hq, d, float(d) ** -0.5, hkv, 0, logit_cap=logit_cap)
    if k_scale is not None:
        layer.k_scale, layer.v_scale = torch.tensor(k_scale), torch.tensor(v_scale)
        layer.k_scale_float, layer.v_scale_float = float(k_scale), float(v_scale)
    return layer


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("scales", [None, (0.5, 2.0), (0.0373, 1.7)])
def test_kv_write_fp8_bit_exact(dtype, scales):
    from sglang_security_b200.mem_cache.memory_pool import MHATokenToKVPool

    g = torch.Generator().manual_seed(7)
    t, hkv, d, size = 37, 4, 128, 300
    k = torch.randn((t, hkv, d), generator=g) * 6
    v = torch.randn((t, hkv, d), generator=g) * 0.01          # exercises e4m3 subnormals
    k[0, 0, :8] = torch.tensor([448., 449., 464., 465., 480., -1000., 0., -0.])  # overflow -> NaN
    k, v = k.to(dtype), v.to(dtype)
    loc = (torch.randperm(size, generator=g)[:t] + 1).to(torch.int64)
    ks, vs = scales if scales else (None, None)
    ref = OracleMHATokenToKVPool(size, 1, FP8, hkv, d, 1)
    ref.set_kv_buffer(SimpleNamespace(layer_id=0), loc, k.clone(), v.clone(),
                      None if ks is None else torch.tensor(ks), None if vs is None else torch.tensor(vs))
    pool = MHATokenToKVPool(size, 1, FP8, hkv, d, 1, "cuda")
    # strided source rows (slices of a fused qkv activation)
    qkv = torch.cat([k.reshape(t, -1), v.reshape(t, -1)], dim=1).cuda()
    kd, vd = qkv.split([hkv * d, hkv * d], dim=-1)
    pool.set_kv_buffer(SimpleNamespace(layer_id=0), loc.cuda(), kd, vd, ks, vs)
    torch.cuda.synchronize()
    assert torch.equal(pool.k_buffer[0].cpu(), ref.k_buffer[0])
    assert torch.equal(pool.v_buffer[0].cpu(), ref.v_buffer[0])
    assert torch.equal(qkv.cpu()[:, :hkv * d], k.reshape(t, -1))  # source left untouched


def _run_decode(case, k_scale, v_scale, num_kv_splits, logit_cap=0.0, fused_tail=None, fp8_out=False,
                tail_in_kernel=None):
    """-> (o_gpu, o_oracle, k bytes gpu, k bytes oracle)"""
    import harness
    from sglang_security_b200.layers.attention.b200_backend import B200AttnBackend
    from sglang_security_b200.model_executor.forward_batch_info import ForwardBatch, ForwardMode

    layer = _layer(case.hq, case.hkv, case.d, k_scale, v_scale, logit_cap)
    ks = 1.0 if k_scale is None else k_scale
    vs = 1.0 if v_scale is None else v_scale
    # oracle: quantise the resident cache the same way new tokens are (cast, no NaNs at N(0,1))
    ref = OracleMHATokenToKVPool(case.pool_size, 1, FP8, case.hkv, case.d, 1)
    ref.k_buffer[0].copy_((case.k_cache.float() / ks).to(case.dtype).to(FP8).view(torch.uint8))
    ref.v_buffer[0].copy_((case.v_cache.float() / vs).to(case.dtype).to(FP8).view(torch.uint8))
    ref.set_kv_buffer(layer, case.out_cache_loc, case.k_new.clone(), case.v_new.clone(),
                      None if k_scale is None else torch.tensor(k_scale),
                      None if v_scale is None else torch.tensor(v_scale))
    o_ref = OA.decode_attention_oracle(
        case.q.view(-1, case.hq, case.d), ref.get_key_buffer(0), ref.get_value_buffer(0),
        case.req_to_token, case.req_pool_indices, case.seq_lens, layer.scaling, logit_cap, ks, vs)
    # device
    fp8_case = SimpleNamespace(**vars(case))
    fp8_case.dtype = FP8
    fp8_case.k_cache = torch.zeros_like(case.k_cache, dtype=torch.uint8)
    fp8_case.v_cache = torch.zeros_like(case.v_cache, dtype=torch.uint8)
    mr = harness.make_model_runner(fp8_case, "cuda", num_kv_splits=num_kv_splits)
    pool = mr.token_to_kv_pool
    init_k, init_v = ref.k_buffer[0].clone(), ref.v_buffer[0].clone()
    init_k[case.out_cache_loc] = 0  # the new rows must come from the device-side write
    init_v[case.out_cache_loc] = 0
    pool.k_buffer[0].copy_(init_k)
    pool.v_buffer[0].copy_(init_v)
    backend = B200AttnBackend(mr)
    if fused_tail is not None:
        backend.fused_tail = fused_tail
    if tail_in_kernel is not None:
        backend.tail_in_kernel_max_bs = 1 << 30 if tail_in_kernel else 0
    t = case.q.shape[0]
    qkv = torch.cat([case.q, case.k_new.reshape(t, -1), case.v_new.reshape(t, -1)], dim=1).cuda()
    q, k, v = qkv.split([case.hq * case.d, case.hkv * case.d, case.hkv * case.d], dim=-1)
    fb = ForwardBatch(forward_mode=ForwardMode.DECODE, batch_size=case.bs,
                      req_pool_indices=case.req_pool_indices.cuda(), seq_lens=case.seq_lens.cuda(),
                      out_cache_loc=case.out_cache_loc.cuda(), seq_lens_sum=int(case.seq_lens.sum()),
                      seq_lens_cpu=case.seq_lens, req_to_token_pool=mr.req_to_token_pool,
                      token_to_kv_pool=pool, attn_backend=backend)
    backend.init_forward_metadata(fb)
    if fp8_out:
        for _ in range(2):
            o = layer(q, k, v, fb, save_kv_cache=False if _ else True, fp8_out=True)
        torch.cuda.synchronize()
        return (o.q.cpu(), o.scale.cpu()), o_ref.reshape(t, -1), pool.k_buffer[0].cpu(), ref.k_buffer[0]
    o = layer(q, k, v, fb)
    torch.cuda.synchronize()
    return o.cpu(), o_ref.reshape(t, -1), pool.k_buffer[0].cpu(), ref.k_buffer[0]


@pytest.mark.parametrize("hq,hkv,d", [(32, 8, 128), (14, 2, 64), (16, 1, 128), (8, 8, 64)])
@pytest.mark.parametrize("splits", [1, 3, 0])
@pytest.mark.parametrize("scales", [None, (0.25, 1.5)])
def test_decode_fp8kv_matches_oracle(hq, hkv, d, splits, scales):
    import harness

    case = scenarios.make_attention_case(seed=hq * 3 + d + splits, hq=hq, hkv=hkv, d=d,
                                         seq_lens=[1, 31, 32, 33, 95, 96, 97, 700, 1025],
                                         shared_prefix=20)
    ks, vs = scales if scales else (None, None)
    o, o_ref, kb, kb_ref = _run_decode(case, ks, vs, splits)
    assert torch.equal(kb, kb_ref), "fp8 KV write not bit-exact"
    st = harness.error_stats(o, o_ref)
    assert st["max_abs"] <= 3e-2 and st["rel_l2"] <= 1e-2, st


def test_decode_fp8kv_f16_q_and_logit_cap():
    import harness

    case = scenarios.make_attention_case(seed=11, hq=16, hkv=4, d=128, seq_lens=[5, 260, 1500],
                                         dtype=torch.float16)
    o, o_ref, kb, kb_ref = _run_decode(case, 0.5, 0.5, 2, logit_cap=30.0)
    assert torch.equal(kb, kb_ref)
    st = harness.error_stats(o, o_ref)
    assert st["max_abs"] <= 3e-2 and st["rel_l2"] <= 1e-2, st


@pytest.mark.parametrize("hq,hkv,d", [(32, 8, 128), (14, 2, 64)])
@pytest.mark.parametrize("scales", [None, (0.5, 1.25)])
def test_extend_fp8kv_matches_oracle(hq, hkv, d, scales):
    """Extend (cached radix prefix + new tokens) over the e4m3 pool: new K/V quantised by the
    write kernel, attention over the exact 16-bit copy of the rows (b200_kv_dequant_gather)."""
    import harness
    from sglang_security_b200.layers.attention.b200_backend import B200AttnBackend
    from sglang_security_b200.model_executor.forward_batch_info import ForwardBatch, ForwardMode

    case = scenarios.make_attention_case(seed=hq + d, hq=hq, hkv=hkv, d=d,
                                         seq_lens=[40, 300, 129, 700], extend_lens=[40, 44, 1, 200],
                                         shared_prefix=64)
    ks, vs = scales if scales else (None, None)
    fks, fvs = (1.0 if ks is None else ks), (1.0 if vs is None else vs)
    layer = _layer(hq, hkv, d, ks, vs)
    ref = OracleMHATokenToKVPool(case.pool_size, 1, FP8, hkv, d, 1)
    ref.k_buffer[0].copy_((case.k_cache.float() / fks).to(case.dtype).to(FP8).view(torch.uint8))
    ref.v_buffer[0].copy_((case.v_cache.float() / fvs).to(case.dtype).to(FP8).view(torch.uint8))
    ref.set_kv_buffer(layer, case.out_cache_loc, case.k_new.clone(), case.v_new.clone(),
                      None if ks is None else torch.tensor(ks), None if vs is None else torch.tensor(vs))
    o_ref = OA.extend_attention_oracle(
        case.q.view(-1, hq, d), ref.get_key_buffer(0), ref.get_value_buffer(0), case.req_to_token,
        case.req_pool_indices, case.seq_lens, case.extend_prefix_lens, case.extend_seq_lens,
        layer.scaling, True, 0.0, fks, fvs)
    fp8_case = SimpleNamespace(**vars(case))
    fp8_case.dtype = FP8
    fp8_case.k_cache = torch.zeros_like(case.k_cache, dtype=torch.uint8)
    fp8_case.v_cache = torch.zeros_like(case.v_cache, dtype=torch.uint8)
    mr = harness.make_model_runner(fp8_case, "cuda")
    pool = mr.token_to_kv_pool
    init_k, init_v = ref.k_buffer[0].clone(), ref.v_buffer[0].clone()
    init_k[case.out_cache_loc] = 0
    init_v[case.out_cache_loc] = 0
    pool.k_buffer[0].copy_(init_k)
    pool.v_buffer[0].copy_(init_v)
    backend = B200AttnBackend(mr)
    t = case.q.shape[0]
    qkv = torch.cat([case.q, case.k_new.reshape(t, -1), case.v_new.reshape(t, -1)], dim=1).cuda()
    q, k, v = qkv.split([hq * d, hkv * d, hkv * d], dim=-1)
    fb = ForwardBatch(forward_mode=ForwardMode.EXTEND, batch_size=case.bs,
                      req_pool_indices=case.req_pool_indices.cuda(), seq_lens=case.seq_lens.cuda(),
                      out_cache_loc=case.out_cache_loc.cuda(), seq_lens_sum=int(case.seq_lens.sum()),
                      seq_lens_cpu=case.seq_lens, extend_seq_lens=case.extend_seq_lens.cuda(),
                      extend_prefix_lens=case.extend_prefix_lens.cuda(),
                      extend_seq_lens_cpu=case.extend_seq_lens.tolist(),
                      extend_prefix_lens_cpu=case.extend_prefix_lens.tolist(),
                      extend_num_tokens=t, req_to_token_pool=mr.req_to_token_pool,
                      token_to_kv_pool=pool, attn_backend=backend)
    backend.init_forward_metadata(fb)
    o = layer(q, k, v, fb)
    torch.cuda.synchronize()
    assert torch.equal(pool.k_buffer[0].cpu(), ref.k_buffer[0])
    assert torch.equal(pool.v_buffer[0].cpu(), ref.v_buffer[0])
    st = harness.error_stats(o.cpu(), o_ref.reshape(t, -1))
    assert st["max_abs"] <= 3e-2 and st["rel_l2"] <= 1e-2, st


@pytest.mark.parametrize("splits", [1, 3])
@pytest.mark.parametrize("in_kernel", [True, False])
def test_decode_fp8kv_fused_tail_equals_unfused(splits, in_kernel):
    """FP8-KV decode: fused merge + per-token quantisation tail == split + merge (+ quantiser)."""
    from sglang_security_b200.layers.quantization.fp8_kernel import sglang_per_token_quant_fp8

    case = scenarios.make_attention_case(seed=41 + splits, hq=32, hkv=8, d=128,
                                         seq_lens=[1, 95, 96, 97, 700, 1025], shared_prefix=20)
    o_plain, _, _, _ = _run_decode(case, 0.25, 1.5, splits, fused_tail=False)
    o_fused, _, _, _ = _run_decode(case, 0.25, 1.5, splits, fused_tail=True, tail_in_kernel=in_kernel)
    assert torch.equal(o_plain.view(torch.int16), o_fused.view(torch.int16))
    (qa, sa), _, _, _ = _run_decode(case, 0.25, 1.5, splits, fused_tail=True, fp8_out=True,
                                    tail_in_kernel=in_kernel)
    q_ref, s_ref = sglang_per_token_quant_fp8(o_plain.cuda())
    assert torch.equal(sa, s_ref.cpu())
    assert torch.equal(qa.view(torch.uint8), q_ref.cpu().view(torch.uint8))
