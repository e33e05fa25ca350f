"""CPU tests: the oracle restatement against the golden vectors produced by the UNMODIFIED
reference backend (oracle/gen_golden.py), and against the live reference when it is mounted."""
import os

import pytest
import torch

from oracle import ref_loader, scenarios
from oracle.pools import kv_indices_oracle

import harness

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")
ALL_CASES = {**scenarios.DECODE_CASES, **scenarios.EXTEND_CASES}


def _golden(name):
    return torch.load(os.path.join(GOLDEN, f"attn_{name}.pt"), weights_only=False)


@pytest.mark.parametrize("name", sorted(ALL_CASES))
def test_restatement_matches_reference_golden(name):
    case = scenarios.make_attention_case(**ALL_CASES[name])
    g = _golden(name)
    assert scenarios.digest(case) == g["input_digest"], "synthetic inputs drifted from the fixture"
    o, kb, vb = harness.oracle_run(case)
    # cache-location writes: bit-exact
    assert torch.equal(kb[case.out_cache_loc], g["k_rows"])
    assert torch.equal(vb[case.out_cache_loc], g["v_rows"])
    # attention: reference = SDPA on 16-bit tensors, restatement = fp32 math, one output rounding.
    # Tolerance of the reference's own CPU kernel test (test/srt/cpu/test_decode.py:136-140).
    st = harness.error_stats(o, g["o"])
    assert st["max_abs"] <= 3e-2 and st["rel_l2"] <= 1e-2, st


@pytest.mark.skipif(not ref_loader.available(), reason="reference tree not mounted")
def test_live_reference_agrees_with_golden():
    from oracle import gen_golden

    backend, attention_type = ref_loader.make_reference_backend()
    name = "decode_llama_g4_d128"
    case = scenarios.make_attention_case(**ALL_CASES[name])
    o, _, _ = gen_golden.run_reference(case, backend, attention_type)
    assert torch.equal(o, _golden(name)["o"])


def test_kv_indices_oracle_matches_reference_test_definition():
    # shapes of test/srt/test_create_kvindices.py:18-71 (table scaled down for CPU time)
    g = torch.Generator().manual_seed(0)
    max_batch, max_ctx = 512, 512
    table = torch.arange(max_batch * max_ctx, dtype=torch.int32).reshape(max_batch, max_ctx)
    for batch in (1, 37, 300):
        req = torch.randperm(max_batch, generator=g)[:batch]
        lens = torch.randperm(max_ctx, generator=g)[:batch]
        indptr, flat = kv_indices_oracle(table, req, lens)
        assert indptr[-1].item() == lens.sum().item() == flat.numel()
        for b in range(batch):
            seg = flat[indptr[b]:indptr[b + 1]]
            assert torch.equal(seg, table[req[b], :lens[b]])


@pytest.mark.parametrize("name", sorted(scenarios.DECODE_CASES))
def test_restatement_matches_reference_cpu_kernel(name):
    """Second, independent reference implementation: the reference's own sgl-kernel CPU decode
    kernel compiled from its sources into oracle/_ref (oracle/build_ref.py)."""
    from oracle import ref_kernels

    if not ref_kernels.available():
        pytest.skip("oracle/_ref not built")
    case = scenarios.make_attention_case(**scenarios.DECODE_CASES[name])
    if case.dtype != torch.bfloat16:
        pytest.skip("reference CPU kernel is bf16-only in its tests")
    o_ref, k_ref, v_ref = harness.oracle_run(case)
    kc, vc = case.k_cache.clone(), case.v_cache.clone()
    o = ref_kernels.decode_attention_ref(
        case.q.view(-1, case.hq, case.d), kc, vc, case.k_new, case.v_new, case.out_cache_loc,
        case.req_to_token, case.req_pool_indices, case.seq_lens, case.d ** -0.5)
    assert torch.equal(kc, k_ref) and torch.equal(vc, v_ref)      # its fused KV write
    st = harness.error_stats(o.reshape(o.shape[0], -1), o_ref)
    assert st["max_abs"] <= 3e-2 and st["rel_l2"] <= 1e-2, st


def _triton_golden():
    return torch.load(os.path.join(GOLDEN, "triton_attn.pt"), weights_only=False)["cases"]


def triton_case(g):
    kw = dict(g["kwargs"])
    kw.pop("logit_cap", None)
    kw["dtype"] = getattr(torch, kw["dtype"].split(".")[-1])
    case = scenarios.make_attention_case(**kw)
    assert scenarios.digest(case) == g["input_digest"], "synthetic inputs drifted from the fixture"
    return case


TRITON_CASES = ["decode_llama_g4_d128", "decode_qwen_g7_d64", "decode_g8_d128", "decode_mha_f16",
                "decode_shared_prefix", "extend_llama_g4_d128", "extend_qwen_g7_d64",
                "extend_shared_prefix", "decode_qwen_g7_d64_cap", "extend_llama_g4_d128_cap"]


@pytest.mark.parametrize("name", TRITON_CASES)
def test_restatement_matches_reference_triton_kernels(name):
    """Third reference implementation: the reference's own Triton decode (split-KV stage 1 + 2,
    MHA and grouped kernels) and extend kernels, run through Triton's CPU interpreter in float16
    (oracle/gen_golden_triton_attn.py).  Both sides do fp32 math with one f16 output rounding, so
    the bound is a few f16 ulps, far inside the 3e-2 / 1e-2 parity tolerance."""
    g = _triton_golden()[name]
    case = triton_case(g)
    o, _, _ = harness.oracle_run(case, logit_cap=g["logit_cap"])
    st = harness.error_stats(o, g["o"])
    assert st["max_abs"] <= 4e-3 and st["rel_l2"] <= 2e-3, st


def test_triton_golden_covers_every_case():
    g = _triton_golden()
    assert sorted(g) == sorted(TRITON_CASES)
    # logit capping changed the result of the capped cases (the fixture really exercises it)
    assert not torch.equal(g["decode_qwen_g7_d64_cap"]["o"], g["decode_qwen_g7_d64"]["o"])
    assert not torch.equal(g["extend_llama_g4_d128_cap"]["o"], g["extend_llama_g4_d128"]["o"])
