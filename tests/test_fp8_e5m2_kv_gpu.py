"""GPU parity of the float8_e5m2 KV pool (`--kv-cache-dtype fp8_e5m2`, model_runner.py:845-846):
the same kernels as the e4m3 pool with the byte conversions swapped (B200_KV_FP8_E5M2 flag).
Writers: bit-exact with torch's div_ -> .to(float8_e5m2) -> index_put (memory_pool.py:390-412,
restated in oracle/pools.py) including the overflow-to-Inf and NaN codes; attention: the e4m3 cases of
test_fp8_kv_gpu.py re-run over an e5m2 pool (same oracle, same tolerances)."""
from types import SimpleNamespace

import pytest
import torch

import test_fp8_kv_gpu as T
from oracle import scenarios
from oracle.pools import OracleMHATokenToKVPool

pytestmark = pytest.mark.gpu
E5M2 = torch.float8_e5m2


@pytest.fixture(autouse=True)
def _e5m2_pool(monkeypatch):
    monkeypatch.setattr(T, "FP8", E5M2)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("scales", [None, (0.5, 2.0), (0.0373, 1.7)])
def test_kv_write_e5m2_bit_exact(dtype, scales):
    from sglang_security_b200.mem_cache.memory_pool import MHATokenToKVPool

    g = torch.Generator().manual_seed(9)
    t, hkv, d, size = 37, 4, 128, 300
    k = torch.randn((t, hkv, d), generator=g) * 6
    v = torch.randn((t, hkv, d), generator=g) * 4e-5          # e5m2 subnormals (< 6.1e-5) and zeros
    edge = [57344., 59000., 61184., 61440., 63488., 65504., 1e5, -1e5, float("inf"), float("-inf"),
            float("nan"), 0., -0., 7.6e-6, 7.7e-6, -2.3e-5]
    k[0, 0, :len(edge)] = torch.tensor(edge)
    k[1, 0, :len(edge)] = torch.tensor(edge) * 0.5
    k, v = k.to(dtype), v.to(dtype)
    loc = (torch.randperm(size, generator=g)[:t] + 1).to(torch.int64)
    ks, vs = scales if scales else (None, None)
    ref = OracleMHATokenToKVPool(size, 1, E5M2, hkv, d, 1)
    ref.set_kv_buffer(SimpleNamespace(layer_id=0), loc, k.clone(), v.clone(),
                      None if ks is None else torch.tensor(ks), None if vs is None else torch.tensor(vs))
    pool = MHATokenToKVPool(size, 1, E5M2, hkv, d, 1, "cuda")
    qkv = torch.cat([k.reshape(t, -1), v.reshape(t, -1)], dim=1).cuda()
    kd, vd = qkv.split([hkv * d, hkv * d], dim=-1)
    pool.set_kv_buffer(SimpleNamespace(layer_id=0), loc.cuda(), kd, vd, ks, vs)
    torch.cuda.synchronize()
    got_k, got_v = pool.k_buffer[0].cpu(), pool.v_buffer[0].cpu()
    # the sign of a NaN that went through div_ is the platform's (the GPU returns the canonical
    # positive NaN, the CPU keeps the operand's): NaN codes are compared without their sign bit here
    # and exactly against torch's own CUDA cast below
    want_k = ref.k_buffer[0].clone()
    nan = (want_k & 0x7f) == 0x7f
    assert torch.equal(nan, (got_k & 0x7f) == 0x7f)
    want_k[nan] = got_k[nan]
    bad = (got_k != want_k).nonzero()
    assert bad.numel() == 0, (bad[:8].tolist(), got_k[tuple(bad[0])].item(), want_k[tuple(bad[0])].item())
    assert torch.equal(got_v, ref.v_buffer[0])
    # the same rows through torch's own CUDA cast (the reference's device path)
    dev = (kd.clone().div_(ks) if ks is not None else kd).to(E5M2).view(torch.uint8).cpu()
    assert torch.equal(got_k[loc].reshape(t, -1), dev)


@pytest.mark.parametrize("hq,hkv,d", [(32, 8, 128), (14, 2, 64), (16, 1, 128), (8, 8, 64)])
@pytest.mark.parametrize("splits", [1, 3, 0])
@pytest.mark.parametrize("scales", [None, (0.25, 1.5)])
def test_decode_e5m2kv_matches_oracle(hq, hkv, d, splits, scales):
    T.test_decode_fp8kv_matches_oracle(hq, hkv, d, splits, scales)


def test_decode_e5m2kv_f16_q_and_logit_cap():
    T.test_decode_fp8kv_f16_q_and_logit_cap()


@pytest.mark.parametrize("hq,hkv,d", [(32, 8, 128), (14, 2, 64)])
@pytest.mark.parametrize("scales", [None, (0.5, 1.25)])
def test_extend_e5m2kv_matches_oracle(hq, hkv, d, scales):
    T.test_extend_fp8kv_matches_oracle(hq, hkv, d, scales)


@pytest.mark.parametrize("splits", [1, 3])
def test_decode_e5m2kv_fused_tail_equals_unfused(splits):
    """The ticket buffer is ignored for an e5m2 pool (no in-kernel tail instance): both settings
    must give the bits of split + merge (+ quantiser)."""
    T.test_decode_fp8kv_fused_tail_equals_unfused(splits, True)
    T.test_decode_fp8kv_fused_tail_equals_unfused(splits, False)


@pytest.mark.parametrize("hq,hkv,d,rot,neox", [(32, 8, 128, 128, True), (14, 2, 64, 64, False)])
@pytest.mark.parametrize("scales", [None, (0.37, 2.5)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_rope_fused_e5m2_kv_write_matches_unfused_bytes(hq, hkv, d, rot, neox, scales, dtype):
    from sglang_security_b200.layers.rotary_embedding import RotaryEmbedding
    from sglang_security_b200.mem_cache.memory_pool import MHATokenToKVPool

    g = torch.Generator().manual_seed(hq + d + (0 if scales is None else 5))
    t, max_pos, pool_size = 41, 2048, 150
    qkv = (torch.randn(t, (hq + 2 * hkv) * d, generator=g) * 3).to(dtype)
    qkv[3, hq * d + 5] = 65000.0 if dtype == torch.float16 else 1e6   # beyond e5m2 range in k -> Inf
    qkv[7, (hq + hkv) * d + 9] = -62000.0                             # and in v
    pos = torch.randint(0, max_pos, (t,), generator=g)
    loc = (torch.randperm(pool_size, generator=g)[:t] + 1).to(torch.int64).cuda()
    rope = RotaryEmbedding(d, rot, max_pos, 10000.0, neox, dtype, native_rounding=True).cuda()
    layer = SimpleNamespace(layer_id=0)
    if scales is not None:
        layer.k_scale_float, layer.v_scale_float = scales

    def run(fused):
        pool = MHATokenToKVPool(pool_size, 1, E5M2, hkv, d, 1, "cuda")
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
    assert (kba[0] == 0x11).all() and (vba[0] == 0x22).all()
