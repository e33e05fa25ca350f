"""GPU parity of the extend (prefill with cached radix prefix) path against the oracle and the
reference goldens (torch_native forward_extend run on CPU)."""
import os

import pytest
import torch

from oracle import scenarios

import harness

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")
# extend tolerance of the reference's own tests: allclose rtol = atol = 1e-2
# (test/srt/test_triton_attention_kernels.py:171-172, test/srt/cpu/test_extend.py:177)
MAX_ABS, REL_L2 = 3e-2, 1e-2


def _check(case, golden=None, **kw):
    o_ref, k_ref, v_ref = harness.oracle_run(case)
    o, kb, vb = harness.b200_run(case, **kw)
    assert torch.equal(kb.view(torch.int16), k_ref.view(torch.int16))
    assert torch.equal(vb.view(torch.int16), v_ref.view(torch.int16))
    st = harness.error_stats(o, o_ref)
    assert st["max_abs"] <= MAX_ABS and st["rel_l2"] <= REL_L2, st
    if golden is not None:
        st = harness.error_stats(o, golden["o"])
        assert st["max_abs"] <= MAX_ABS and st["rel_l2"] <= REL_L2, ("vs reference golden", st)
    return st


@pytest.mark.parametrize("name", sorted(scenarios.EXTEND_CASES))
def test_extend_matches_oracle_and_golden(name):
    case = scenarios.make_attention_case(**scenarios.EXTEND_CASES[name])
    g = torch.load(os.path.join(GOLDEN, f"attn_{name}.pt"), weights_only=False)
    _check(case, golden=g)


@pytest.mark.parametrize("hq,hkv,d", [(32, 8, 128), (14, 2, 64), (8, 1, 128), (4, 4, 128),
                                      (40, 8, 128), (6, 2, 64)])
def test_extend_geometries_ragged(hq, hkv, d):
    # prefix / extend lengths around the 32/64/128 tile boundaries, incl. no prefix, 1-token
    # extends (MIXED mode decodes) and prefix >> extend
    seq = [1, 64, 65, 200, 130, 513, 33, 700]
    ext = [1, 64, 1, 137, 130, 1, 32, 129]
    case = scenarios.make_attention_case(seed=hq * 10 + d, hq=hq, hkv=hkv, d=d, seq_lens=seq,
                                         extend_lens=ext)
    _check(case)


def test_extend_shared_radix_prefix_and_f16():
    case = scenarios.make_attention_case(seed=77, hq=8, hkv=2, d=128, seq_lens=[528] * 4 + [600],
                                         extend_lens=[16, 16, 16, 16, 88], shared_prefix=512,
                                         dtype=torch.float16)
    _check(case)


def test_extend_long_prefix_chunked_prefill_shape():
    # one chunk of a chunked prefill: 256 new tokens after 3000 cached ones (lazy-rescale path
    # exercised by the N(0,1) score range), plus a pure prefill of 300
    case = scenarios.make_attention_case(seed=78, hq=8, hkv=2, d=128, seq_lens=[3256, 300],
                                         extend_lens=[256, 300])
    _check(case)


def test_extend_large_magnitude_scores_rescale():
    # std 3 inputs -> logits with a wide range: forces several O rescales per row
    case = scenarios.make_attention_case(seed=79, hq=4, hkv=1, d=128, seq_lens=[400, 150],
                                         extend_lens=[200, 150], std=3.0)
    o_ref, _, _ = harness.oracle_run(case)
    o, _, _ = harness.b200_run(case)
    st = harness.error_stats(o, o_ref)
    assert st["rel_l2"] <= 2e-2, st


def test_extend_int32_index_tensors():
    """int32 req_pool_indices / seq_lens (cuda_graph_runner.py:250-253 convention): converted copies
    are held in the metadata for the whole forward, never as temporaries."""
    case = scenarios.make_attention_case(seed=80, hq=8, hkv=2, d=128, seq_lens=[700, 64, 65, 300],
                                         extend_lens=[129, 64, 1, 300])
    _check(case, index_dtype=torch.int32)


def test_extend_8k_context_llama_heads():
    """cfg-3/5 chunk shape at the size the perf numbers are quoted on: 512 new tokens after a 7680-
    token cached prefix (8k context), Llama-3.1-8B heads, and a radix-hit-0.9 request."""
    case = scenarios.make_attention_case(seed=81, hq=32, hkv=8, d=128, seq_lens=[8192, 2048],
                                         extend_lens=[512, 205])
    _check(case)


def test_extend_tp8_rank_geometry():
    """Llama-3.1-70B at TP=8: a rank holds Hq 8 / Hkv 1 (one GQA group fills the whole M tile)."""
    case = scenarios.make_attention_case(seed=82, hq=8, hkv=1, d=128, seq_lens=[8192, 1024, 130],
                                         extend_lens=[1024, 1024, 1])
    _check(case)
