"""CPU: the reference's fused-shard weight-loader contract (python/sglang/srt/layers/linear.py:
523-750 MergedColumnParallelLinear, :937-1128 QKVParallelLinear; parameter.py:88-195,340-375):
`loaded_shard_id`, per-rank windows, kv-head replication, int4 packing, block-scale grids and
per-shard scalar scales.  Every case loads the SAME full checkpoint on every simulated rank and
compares the rank's parameter with a hand-computed slice; GPTQ row-parallel g_idx handling
 and the Llama-3.1 rope_scaling table (pinned to the reference's get_rope output)."""
import os

import pytest
import torch

from sglang_security_b200.layers.linear import (MergedColumnParallelLinear, QKVParallelLinear,
                                                RowParallelLinear)
from sglang_security_b200.layers.quantization.awq import AWQConfig
from sglang_security_b200.layers.quantization.fp8 import Fp8Config
from sglang_security_b200.layers.quantization.gptq import GPTQConfig

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


@pytest.mark.parametrize("tp", [1, 2, 4])
def test_merged_column_shard_ids_and_fused_checkpoint(tp):
    hidden, inter = 32, 48
    g = torch.Generator().manual_seed(0)
    gate, up = torch.randn(inter, hidden, generator=g), torch.randn(inter, hidden, generator=g)
    shard = inter // tp
    for rank in range(tp):
        want = torch.cat([gate[rank * shard:(rank + 1) * shard], up[rank * shard:(rank + 1) * shard]])
        a = MergedColumnParallelLinear(hidden, [inter, inter], bias=False, params_dtype=torch.float32,
                                       tp_rank=rank, tp_size=tp)
        a.weight.weight_loader(a.weight, gate, 0)
        a.weight.weight_loader(a.weight, up, 1)
        assert torch.equal(a.weight.data, want)
        b = MergedColumnParallelLinear(hidden, [inter, inter], bias=False, params_dtype=torch.float32,
                                       tp_rank=rank, tp_size=tp)
        b.weight.weight_loader(b.weight, torch.cat([gate, up]))      # fused on disk, no shard id
        assert torch.equal(b.weight.data, want)


def test_merged_column_rejects_bad_shard():
    a = MergedColumnParallelLinear(8, [16, 16], bias=False, params_dtype=torch.float32, tp_rank=0, tp_size=2)
    with pytest.raises(ValueError):
        a.weight.weight_loader(a.weight, torch.zeros(16, 8), 2)
    with pytest.raises(ValueError):
        a.weight.weight_loader(a.weight, torch.zeros(4, 8), 0)        # narrower than this rank's window


@pytest.mark.parametrize("tp,heads,kv_heads", [(1, 8, 2), (2, 8, 2), (4, 8, 2), (8, 8, 2)])
def test_qkv_loader_with_replicated_kv_heads(tp, heads, kv_heads):
    hidden, d = 24, 4
    g = torch.Generator().manual_seed(1)
    q = torch.randn(heads * d, hidden, generator=g)
    k = torch.randn(kv_heads * d, hidden, generator=g)
    v = torch.randn(kv_heads * d, hidden, generator=g)
    bq, bk, bv = (torch.randn(t.shape[0], generator=g) for t in (q, k, v))
    for rank in range(tp):
        layer = QKVParallelLinear(hidden, d, heads, kv_heads, bias=True, params_dtype=torch.float32,
                                  tp_rank=rank, tp_size=tp)
        hq, hkv = heads // tp, max(1, kv_heads // tp)
        kv_rank = rank // layer.num_kv_head_replicas
        assert layer.num_kv_head_replicas == max(1, tp // kv_heads)
        rows = lambda t, n, r: t[r * n * d:(r + 1) * n * d]
        want_w = torch.cat([rows(q, hq, rank), rows(k, hkv, kv_rank), rows(v, hkv, kv_rank)])
        want_b = torch.cat([rows(bq, hq, rank), rows(bk, hkv, kv_rank), rows(bv, hkv, kv_rank)])
        for sid, w, b in (("q", q, bq), ("k", k, bk), ("v", v, bv)):
            layer.weight.weight_loader(layer.weight, w, sid)
            layer.bias.weight_loader(layer.bias, b, sid)
        assert torch.equal(layer.weight.data, want_w) and torch.equal(layer.bias.data, want_b)
        fused = QKVParallelLinear(hidden, d, heads, kv_heads, bias=False, params_dtype=torch.float32,
                                  tp_rank=rank, tp_size=tp)
        fused.weight.weight_loader(fused.weight, torch.cat([q, k, v]))
        assert torch.equal(fused.weight.data, want_w)
    with pytest.raises(ValueError):
        layer.weight.weight_loader(layer.weight, q, "x")


def test_awq_packed_parameters_use_packed_windows():
    """qweight / qzeros are packed 8-to-1 along the output dim (awq.py:139-165): windows shrink by
    the pack factor (parameter.py:340-375 adjust_shard_indexes_for_packing); scales do not."""
    k, inter, group, tp = 128, 64, 64, 2
    cfg = AWQConfig(4, group, True)
    g = torch.Generator().manual_seed(2)
    ck = {}
    for name in ("gate", "up"):
        ck[name] = dict(qweight=torch.randint(-2 ** 31, 2 ** 31 - 1, (k, inter // 8), generator=g, dtype=torch.int32),
                        qzeros=torch.randint(-2 ** 31, 2 ** 31 - 1, (k // group, inter // 8), generator=g, dtype=torch.int32),
                        scales=torch.rand(k // group, inter, generator=g).to(torch.float16))
    for rank in range(tp):
        layer = MergedColumnParallelLinear(k, [inter, inter], bias=False, params_dtype=torch.float16,
                                           quant_config=cfg, tp_rank=rank, tp_size=tp)
        for sid, name in enumerate(("gate", "up")):
            for pname, t in ck[name].items():
                p = getattr(layer, pname)
                p.weight_loader(p, t, sid)
        for pname in ("qweight", "qzeros", "scales"):
            cols = ck["gate"][pname].shape[1] // tp
            want = torch.cat([ck["gate"][pname][:, rank * cols:(rank + 1) * cols],
                              ck["up"][pname][:, rank * cols:(rank + 1) * cols]], dim=1)
            assert torch.equal(getattr(layer, pname).data, want), pname


def test_fp8_per_tensor_scales_fill_their_shard_slot():
    """Fp8 serialized checkpoints carry one scalar weight_scale per logical projection
    (fp8.py:236-262; linear.py:78-94 adjust_scalar_to_fused_array)."""
    cfg = Fp8Config(is_checkpoint_fp8_serialized=True, activation_scheme="static")
    layer = QKVParallelLinear(32, 4, 8, 2, bias=False, params_dtype=torch.bfloat16, quant_config=cfg,
                              tp_rank=1, tp_size=2)
    for sid, val in (("q", 0.5), ("k", 0.25), ("v", 2.0)):
        layer.weight_scale.weight_loader(layer.weight_scale, torch.tensor(val), sid)
        layer.input_scale.weight_loader(layer.input_scale, torch.tensor(val * 3), sid)
    assert layer.weight_scale.data.tolist() == [0.5, 0.25, 2.0]
    assert layer.input_scale.data.tolist() == [1.5, 0.75, 6.0]
    layer.weight_scale.weight_loader(layer.weight_scale, torch.tensor(0.125))   # fused on disk
    assert layer.weight_scale.data.tolist() == [0.125] * 3
    w = torch.randn(8 * 4, 32).to(torch.float8_e4m3fn)
    layer.weight.weight_loader(layer.weight, w, "q")
    assert torch.equal(layer.weight.data[:16].view(torch.uint8), w[16:32].view(torch.uint8))


def test_fp8_block_scale_grid_windows():
    """128x128 block scales: a logical projection of R rows owns ceil(R / block_n) grid rows
    (linear.py:731-741)."""
    cfg = Fp8Config(is_checkpoint_fp8_serialized=True, activation_scheme="dynamic", weight_block_size=[128, 128])
    hidden, inter, tp = 256, 512, 2
    g = torch.Generator().manual_seed(3)
    sg, su = torch.rand(inter // 128, hidden // 128, generator=g), torch.rand(inter // 128, hidden // 128, generator=g)
    for rank in range(tp):
        layer = MergedColumnParallelLinear(hidden, [inter, inter], bias=False, params_dtype=torch.bfloat16,
                                           quant_config=cfg, tp_rank=rank, tp_size=tp)
        p = layer.weight_scale_inv
        p.weight_loader(p, sg, 0)
        p.weight_loader(p, su, 1)
        r = inter // 128 // tp
        assert torch.equal(p.data, torch.cat([sg[rank * r:(rank + 1) * r], su[rank * r:(rank + 1) * r]]))


def test_gptq_row_parallel_ignores_global_g_idx_without_desc_act():
    """AutoGPTQ stores g_idx = arange(K) // group with GLOBAL group ids even when desc_act=False;
    a K shard of rank > 0 must still be treated as the trivial mapping."""
    k, n, group, tp = 256, 16, 64, 2
    cfg = GPTQConfig(4, group, False)
    for rank in range(tp):
        layer = RowParallelLinear(k, n, bias=False, params_dtype=torch.float16, quant_config=cfg,
                                  tp_rank=rank, tp_size=tp)
        g_idx = torch.arange(k, dtype=torch.int32) // group
        layer.g_idx.weight_loader(layer.g_idx, g_idx)
        assert layer.g_idx.shape == (k // tp,)
        method = layer.quant_method
        import sglang_security_b200.layers.quantization.gptq as G
        calls = {}
        orig = (G.repack_gptq, G.unpack_zeros, G.gemv_stream_repack)
        G.repack_gptq = lambda qw, perm: calls.setdefault("perm", perm) or qw
        G.unpack_zeros = lambda *a, **k_: None
        G.gemv_stream_repack = lambda qw: None
        try:
            method.process_weights_after_loading(layer)
        finally:
            G.repack_gptq, G.unpack_zeros, G.gemv_stream_repack = orig
        assert calls["perm"] is None and layer.b200_perm is None
    with pytest.raises(NotImplementedError):
        RowParallelLinear(k, n, bias=False, params_dtype=torch.float16,
                          quant_config=GPTQConfig(4, group, True), tp_rank=1, tp_size=2)


def test_llama3_rope_table_matches_reference_get_rope():
    """tests/golden/elementwise_rope_llama3_d128.pt = cos_sin_cache of the UNMODIFIED reference
    get_rope(..., rope_scaling=llama3) (oracle/gen_golden_elementwise.py)."""
    from sglang_security_b200.layers.rotary_embedding import Llama3RotaryEmbedding, get_rope

    gold = torch.load(os.path.join(GOLDEN, "elementwise_rope_llama3_d128.pt"))
    rope = get_rope(gold["head_size"], gold["rotary_dim"], gold["max_pos"], gold["base"], True,
                    gold["rope_scaling"], torch.bfloat16)
    assert isinstance(rope, Llama3RotaryEmbedding)
    assert torch.equal(rope.cos_sin_cache, gold["cos_sin_cache"])
    plain = get_rope(gold["head_size"], gold["rotary_dim"], gold["max_pos"], gold["base"], True, None)
    assert not torch.equal(plain.cos_sin_cache, gold["cos_sin_cache"])
    with pytest.raises(NotImplementedError):
        get_rope(128, 128, 2048, 10000.0, True, {"rope_type": "yarn", "factor": 4.0})
