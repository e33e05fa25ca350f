"""GPU parity of the fused steps either side of the GEMMs / attention (b200_rmsnorm,
b200_silu_and_mul, b200_rope_kv_write) against the CPU oracle (oracle/elementwise.py, pinned to the
reference's forward_native by tests/golden/elementwise_*.pt) and against the unfused chain of our
own kernels (bit-exact)."""
import os

import pytest
import torch

from oracle import elementwise as E
from oracle import fp8 as OF

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")
FP8 = torch.float8_e4m3fn


def _ulp_close(a, b, ulps):
    """|a-b| <= ulps * ulp(|b|) for 16-bit tensors (8 significand bits bf16 / 11 f16)."""
    bits = 8 if a.dtype == torch.bfloat16 else 11
    af, bf = a.float(), b.float()
    ulp = torch.exp2(torch.floor(torch.log2(bf.abs().clamp_min(1e-30))) - (bits - 1))
    bad = (af - bf).abs() > ulps * ulp
    return int(bad.sum())


def _load(name):
    return torch.load(os.path.join(GOLDEN, f"elementwise_{name}.pt"))


@pytest.mark.parametrize("name", ["rms_896_bf16", "rms_1024_f16"])
def test_rmsnorm_golden(name):
    from sglang_security_b200.layers.layernorm import RMSNorm

    c = _load(name)
    k = c["x"].shape[1]
    norm = RMSNorm(k, c["eps"]).cuda()
    norm.weight.data = c["weight"].cuda()
    y = norm(c["x"].cuda())
    assert _ulp_close(y.cpu(), c["y"], 1) == 0
    r = c["residual"].cuda().clone()
    y2, r2 = norm(c["x"].cuda(), r)
    assert r2.data_ptr() == r.data_ptr()
    assert torch.equal(r2.cpu(), c["residual_out"])  # fp32 add, one rounding: bit-exact
    assert _ulp_close(y2.cpu(), c["y_res"], 1) == 0


@pytest.mark.parametrize("m,k", [(1, 4096), (64, 4096), (7, 8192), (3, 16384), (300, 5120), (2, 896)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_rmsnorm_seeded_and_fused_quant(m, k, dtype):
    from sglang_security_b200.layers.layernorm import RMSNorm
    from sglang_security_b200.layers.quantization.fp8_kernel import (per_token_group_quant_fp8,
                                                                    sglang_per_token_quant_fp8)

    g = torch.Generator().manual_seed(m * 13 + k)
    x = torch.randn(m, k, generator=g).to(dtype)
    res = torch.randn(m, k, generator=g).to(dtype)
    w = (1 + 0.1 * torch.randn(k, generator=g)).to(dtype)
    y_ref, r_ref = E.rmsnorm_oracle(x, w, 1e-5, res)
    norm = RMSNorm(k, 1e-5).cuda()
    norm.weight.data = w.cuda()
    r = res.cuda().clone()
    y, _ = norm(x.cuda(), r)
    assert torch.equal(r.cpu(), r_ref)
    # reduction order differs from torch's CPU mean: allow 1 ulp on a handful of elements
    assert _ulp_close(y.cpu(), y_ref, 1) == 0
    # fused norm->quant emits exactly the bytes of the unfused chain of our kernels
    q_ref, s_ref = sglang_per_token_quant_fp8(y)
    r = res.cuda().clone()
    q, s = norm.forward_quant_fp8(x.cuda(), r)
    assert torch.equal(r.cpu(), r_ref)
    assert torch.equal(s, s_ref) and torch.equal(q.view(torch.uint8), q_ref.view(torch.uint8))
    # and the oracle's per-token quantiser applied to OUR y gives the same bytes (quantiser parity)
    q_o, s_o = OF.per_token_quant_fp8_oracle(y.cpu())
    assert torch.equal(s.cpu(), s_o) and torch.equal(q.cpu().view(torch.uint8), q_o.view(torch.uint8))
    if k % 128 == 0:
        for col_major in (False, True):
            qg_ref, sg_ref = per_token_group_quant_fp8(y, 128, column_major_scales=col_major)
            qg, sg = norm.forward_quant_fp8(x.cuda(), res.cuda().clone(), group_size=128,
                                            column_major_scales=col_major)
            assert torch.equal(sg, sg_ref)
            assert torch.equal(qg.view(torch.uint8), qg_ref.view(torch.uint8))


@pytest.mark.parametrize("m,d", [(1, 14336), (64, 14336), (5, 256), (33, 4864), (130, 3584),
                                 (3, 4864), (16, 14336), (20, 5120), (2, 16384), (4, 2056),
                                 (3, 27648), (200, 27648)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_silu_and_mul(m, d, dtype):
    from sglang_security_b200.layers.activation import SiluAndMul
    from sglang_security_b200.layers.quantization.fp8_kernel import (per_token_group_quant_fp8,
                                                                    sglang_per_token_quant_fp8)

    g = torch.Generator().manual_seed(m + d)
    x = (torch.randn(m, 2 * d, generator=g) * 2).to(dtype)
    ref = E.silu_and_mul_oracle(x)
    act = SiluAndMul()
    y = act(x.cuda())
    # expf vs the CPU's vectorised exp may differ in the last fp32 bit: <= 1 ulp of the 16-bit result
    assert _ulp_close(y.cpu(), ref, 1) == 0
    assert (y.cpu() != ref).float().mean().item() < 2e-3
    q_ref, s_ref = sglang_per_token_quant_fp8(y)
    q, s = act.forward_quant_fp8(x.cuda())
    assert torch.equal(s, s_ref) and torch.equal(q.view(torch.uint8), q_ref.view(torch.uint8))
    if d % 128 == 0:
        qg_ref, sg_ref = per_token_group_quant_fp8(y, 128)
        qg, sg = act.forward_quant_fp8(x.cuda(), group_size=128)
        assert torch.equal(sg, sg_ref) and torch.equal(qg.view(torch.uint8), qg_ref.view(torch.uint8))


def test_silu_golden():
    from sglang_security_b200.layers.activation import SiluAndMul

    c = _load("silu_256_bf16")
    y = SiluAndMul()(c["x"].cuda())
    assert _ulp_close(y.cpu(), c["y"], 1) == 0


@pytest.mark.parametrize("name", ["rope_neox_d64", "rope_neox_partial_d128", "rope_gptj_d128"])
def test_rope_golden_bit_exact_native_rounding(name):
    from sglang_security_b200.layers.rotary_embedding import RotaryEmbedding

    c = _load(name)
    rope = RotaryEmbedding(c["head_size"], c["rotary_dim"], c["max_pos"], c["base"], c["is_neox"],
                           torch.bfloat16, native_rounding=True).cuda()
    assert torch.equal(rope.cos_sin_cache.cpu(), c["cos_sin_cache"])
    q, k = c["q"].cuda().clone(), c["k"].cuda().clone()
    qo, ko = rope(c["positions"].cuda(), q, k)
    assert qo.data_ptr() == q.data_ptr()  # in place
    assert torch.equal(qo.cpu(), c["q_out"]) and torch.equal(ko.cpu(), c["k_out"])
    # fp32 single-rounding mode (the reference's CUDA kernel semantics): within 1 ulp + exact products
    rope.native_rounding = False
    q, k = c["q"].cuda().clone(), c["k"].cuda().clone()
    rope(c["positions"].cuda(), q, k)
    ref = E.rope_oracle(c["positions"], c["q"].float(), c["k"].float(), c["cos_sin_cache"],
                        c["head_size"], c["rotary_dim"], c["is_neox"])
    assert _ulp_close(q.cpu(), ref[0].to(torch.bfloat16), 1) == 0
    assert _ulp_close(k.cpu(), ref[1].to(torch.bfloat16), 1) == 0


@pytest.mark.parametrize("hq,hkv,d,rot,neox", [(32, 8, 128, 128, True), (14, 2, 64, 64, True),
                                               (8, 8, 128, 64, False), (4, 1, 128, 128, True)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_rope_fused_kv_write_bit_exact(hq, hkv, d, rot, neox, dtype):
    """q/k are strided slices of a fused qkv activation; rotated k and v land in the pool rows
    out_cache_loc exactly as rope followed by set_kv_buffer would put them; other rows untouched."""
    from types import SimpleNamespace

    from sglang_security_b200.layers.rotary_embedding import RotaryEmbedding
    from sglang_security_b200.mem_cache.memory_pool import MHATokenToKVPool

    g = torch.Generator().manual_seed(hq * 7 + d)
    t, max_pos, pool_size = 37, 4096, 200
    qkv = torch.randn(t, (hq + 2 * hkv) * d, generator=g).to(dtype)
    pos = torch.randint(0, max_pos, (t,), generator=g)
    loc = (torch.randperm(pool_size, generator=g)[:t] + 1).to(torch.int64)
    cache = E.cos_sin_cache_oracle(rot, max_pos, 500000.0)
    q_ref, k_ref = E.rope_oracle(pos, qkv[:, :hq * d].contiguous(),
                                 qkv[:, hq * d:(hq + hkv) * d].contiguous(), cache, d, rot, neox)
    rope = RotaryEmbedding(d, rot, max_pos, 500000.0, neox, dtype, native_rounding=True).cuda()
    pool = MHATokenToKVPool(pool_size, 1, dtype, hkv, d, 2, "cuda")
    pool.k_buffer[1].fill_(3.0)
    pool.v_buffer[1].fill_(-2.0)
    dq = qkv.cuda()
    q, k, v = dq.split([hq * d, hkv * d, hkv * d], dim=-1)
    fb = SimpleNamespace(token_to_kv_pool=pool, out_cache_loc=loc.cuda())
    rope.forward_with_kv_write(pos.cuda(), q, k, v, SimpleNamespace(layer_id=1), fb)
    assert torch.equal(q.cpu(), q_ref) and torch.equal(k.cpu(), k_ref)
    kb, vb = pool.k_buffer[1].cpu(), pool.v_buffer[1].cpu()
    assert torch.equal(kb[loc].reshape(t, -1), k_ref)
    assert torch.equal(vb[loc].reshape(t, -1), qkv[:, (hq + hkv) * d:])
    untouched = torch.ones(pool_size + 1, dtype=torch.bool)
    untouched[loc] = False
    assert (kb[untouched] == 3.0).all() and (vb[untouched] == -2.0).all()
    assert (pool.k_buffer[0] == 0).all()


@pytest.mark.parametrize("hq,hkv,d,rot,neox", [(32, 8, 128, 128, True), (8, 8, 128, 64, False),
                                               (14, 2, 64, 64, True)])
@pytest.mark.parametrize("scales", [None, (0.37, 2.5)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_rope_fused_fp8_kv_write_matches_unfused_bytes(hq, hkv, d, rot, neox, scales, dtype):
    """e4m3 pool: the one-launch RoPE + quantising scatter writes the bytes that RoPE followed by
    set_kv_buffer (b200_kv_write_fp8, itself pinned to torch's div_ -> .to(float8_e4m3fn) in
    test_fp8_kv_gpu.py) writes, including out-of-range values (NaN code) and untouched rows."""
    from types import SimpleNamespace

    from sglang_security_b200.layers.rotary_embedding import RotaryEmbedding
    from sglang_security_b200.mem_cache.memory_pool import MHATokenToKVPool

    g = torch.Generator().manual_seed(hq + d + (0 if scales is None else 5))
    t, max_pos, pool_size = 41, 2048, 150
    qkv = (torch.randn(t, (hq + 2 * hkv) * d, generator=g) * 3).to(dtype)
    qkv[3, hq * d + 5] = 700.0          # beyond e4m3 range in k
    qkv[7, (hq + hkv) * d + 9] = -900.0  # and in v
    pos = torch.randint(0, max_pos, (t,), generator=g)
    loc = (torch.randperm(pool_size, generator=g)[:t] + 1).to(torch.int64).cuda()
    rope = RotaryEmbedding(d, rot, max_pos, 10000.0, neox, dtype, native_rounding=True).cuda()
    layer = SimpleNamespace(layer_id=0)
    if scales is not None:
        layer.k_scale_float, layer.v_scale_float = scales

    def run(fused):
        pool = MHATokenToKVPool(pool_size, 1, torch.float8_e4m3fn, hkv, d, 1, "cuda")
        pool.k_buffer[0].view(torch.uint8).fill_(0x11)
        pool.v_buffer[0].view(torch.uint8).fill_(0x22)
        dq = qkv.cuda().clone()
        q, k, v = dq.split([hq * d, hkv * d, hkv * d], dim=-1)
        if fused:
            fb = SimpleNamespace(token_to_kv_pool=pool, out_cache_loc=loc)
            rope.forward_with_kv_write(pos.cuda(), q, k, v, layer, fb)
        else:
            rope(pos.cuda(), q, k)
            ks, vs = scales if scales is not None else (None, None)
            pool.set_kv_buffer(layer, loc, k.reshape(t, hkv, d), v.reshape(t, hkv, d), ks, vs)
        return (q.clone(), k.clone(), pool.k_buffer[0].view(torch.uint8).clone(),
                pool.v_buffer[0].view(torch.uint8).clone())

    qa, ka, kba, vba = run(True)
    qb, kb, kbb, vbb = run(False)
    assert torch.equal(qa, qb) and torch.equal(ka, kb)
    assert torch.equal(kba, kbb) and torch.equal(vba, vbb)
    assert (kba[loc.cpu()] != 0x11).any() and (kba[0] == 0x11).all() and (vba[0] == 0x22).all()
