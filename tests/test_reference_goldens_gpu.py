"""GPU: the CUDA kernels (through the C-ABI) against fixtures produced by the reference's OWN in-test
torch functions (oracle/gen_golden_quant.py).  Integer / byte results bit-exact; GEMMs within the
reference tests' tolerances (test_fp8_gemm.py:33-35 rtol 0.02 / atol 1; test_block_fp8.py:392-396
relative mean error < 1e-3) and the tighter bound of one output rounding."""
import os

import pytest
import torch

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")
FP8 = torch.float8_e4m3fn


def _load(name):
    return torch.load(os.path.join(GOLD, name + ".pt"), weights_only=False)


def test_awq_dequant_kernel_equals_reference_function():
    from sglang_security_b200.layers.quantization.w4a16_kernel import awq_dequantize

    for c in _load("awq_dequant"):
        out = awq_dequantize(c["qweight"].cuda(), c["scales"].cuda(), c["qzeros"].cuda())
        assert torch.equal(out.cpu().view(torch.int16), c["out"].view(torch.int16))


def test_fp8_quant_kernels_equal_reference_functions():
    from sglang_security_b200.layers.quantization.fp8_kernel import (per_token_group_quant_fp8,
                                                                    sglang_per_token_quant_fp8,
                                                                    static_quant_fp8)

    for c in _load("fp8_quant"):
        q, s = sglang_per_token_quant_fp8(c["x"].cuda())
        assert torch.equal(s.cpu().reshape(-1), c["token_scale"])
        keep = torch.ones(c["x"].shape[0], dtype=torch.bool)
        keep[c["zero_rows"]] = False
        assert torch.equal(q.cpu()[keep].view(torch.uint8), c["q_token"][keep].view(torch.uint8))
        assert (q.cpu()[~keep].float() == 0).all()
        qg, sg = per_token_group_quant_fp8(c["x_group"].cuda(), 128)
        live = c["x_group"].float().reshape(c["x"].shape[0], -1, 128).abs().amax(-1) != 0
        assert torch.equal(sg.cpu()[live], c["s_group"][live])
        live_q = live.repeat_interleave(128, dim=1)
        assert torch.equal(qg.cpu().view(torch.uint8)[live_q], c["q_group"].view(torch.uint8)[live_q])
        # all-zero groups follow the reference KERNELS' fp32 eps clamp (see test_oracle_pinned.py)
        assert (sg.cpu()[~live] == torch.tensor(1e-10) / 448.0).all()
        qs, _ = static_quant_fp8(c["x"].cuda(), c["static_scale"].cuda())
        assert torch.equal(qs.cpu().view(torch.uint8), c["q_static"].view(torch.uint8))


def test_fp8_scaled_mm_kernel_vs_reference_function():
    from sglang_security_b200.layers.quantization.fp8_kernel import fp8_scaled_mm

    for c in _load("fp8_gemm"):
        ref = c["out"]
        bias = None if c["bias"] is None else c["bias"].cuda()
        out = fp8_scaled_mm(c["a"].cuda(), c["b"].cuda().t(), c["scale_a"].cuda(),
                            c["scale_b"].cuda(), ref.dtype, bias)
        torch.testing.assert_close(out.cpu(), ref, rtol=0.02, atol=1)
        err = (out.cpu().float() - ref.float()).abs().max().item()
        assert err <= 2e-2 * max(1.0, ref.float().abs().max().item()), err


def test_fp8_block_mm_kernel_vs_reference_function():
    from sglang_security_b200.layers.quantization.fp8_kernel import fp8_blockwise_scaled_mm

    for c in _load("fp8_block"):
        ref = c["out"]
        out = fp8_blockwise_scaled_mm(c["a"].cuda(), c["b"].cuda(), c["a_s"].cuda(),
                                      c["b_s"].cuda(), ref.dtype)
        rel = (out.cpu().float() - ref.float()).abs().mean() / ref.float().abs().mean()
        assert rel.item() < 1e-3, rel.item()


def test_fp8_pool_write_kernel_equals_reference_method():
    """b200_kv_write_fp8 (e4m3 and e5m2 pools) against the bytes the reference's own
    MHATokenToKVPool.set_kv_buffer wrote on CPU (tests/golden/kv_write_fp8.pt,
    oracle/gen_golden_kv_write.py); untouched rows stay zero."""
    from types import SimpleNamespace

    from sglang_security_b200.mem_cache.memory_pool import MHATokenToKVPool

    for c in _load("kv_write_fp8")["cases"]:
        pool_dtype = getattr(torch, c["pool_dtype"].split(".")[-1])
        t, hkv, d = c["k"].shape
        pool = MHATokenToKVPool(100, 1, pool_dtype, hkv, d, 1, "cuda")
        ks, vs = c["scales"] if c["scales"] else (None, None)
        pool.set_kv_buffer(SimpleNamespace(layer_id=0), c["loc"].cuda(), c["k"].cuda(), c["v"].cuda(), ks, vs)
        torch.cuda.synchronize()
        kb, vb = pool.k_buffer[0].cpu(), pool.v_buffer[0].cpu()
        assert torch.equal(kb[c["loc"]], c["k_rows"]), (c["pool_dtype"], c["dtype"], c["scales"])
        assert torch.equal(vb[c["loc"]], c["v_rows"])
        untouched = torch.ones(kb.shape[0], dtype=torch.bool)
        untouched[c["loc"]] = False
        assert not kb[untouched].any() and not vb[untouched].any()


@pytest.mark.parametrize("splits", [0, 3])
def test_attention_kernels_vs_reference_triton_kernels(splits):
    """B200AttnBackend decode / extend against the outputs of the reference's own Triton kernels
    (tests/golden/triton_attn.pt: decode stage 1 + 2 and extend, run through Triton's CPU
    interpreter by oracle/gen_golden_triton_attn.py), float16, incl. logit capping."""
    import harness
    from oracle import scenarios

    gold = _load("triton_attn")["cases"]
    assert len(gold) == 10
    for name, g in gold.items():
        kw = dict(g["kwargs"])
        kw.pop("logit_cap", None)
        kw["dtype"] = getattr(torch, kw["dtype"].split(".")[-1])
        case = scenarios.make_attention_case(**kw)
        assert scenarios.digest(case) == g["input_digest"]
        o, _, _ = harness.b200_run(case, logit_cap=g["logit_cap"], num_kv_splits=splits)
        st = harness.error_stats(o, g["o"])
        assert st["max_abs"] <= 3e-2 and st["rel_l2"] <= 1e-2, (name, st)
