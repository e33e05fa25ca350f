"""CPU: the B200 linear layers' weight loaders against the parameters the reference's OWN loaders
produce (parameter.py load_*_weight + linear.py weight_loader_v2 of Column / MergedColumn / QKV /
RowParallelLinear, run by oracle/gen_golden_weight_loaders.py -> tests/golden/weight_loaders.pt):
the same full checkpoint is loaded on every simulated TP rank, with shard ids and — where the reference
supports it — fused on disk; unquantised, AWQ (packed output dim), GPTQ (packed input dim), FP8
per-tensor and 128x128 block scales, kv-head replication at tp > kv heads.  Bit-exact."""
import os

import pytest
import torch

from sglang_security_b200.layers.linear import (ColumnParallelLinear, MergedColumnParallelLinear,
                                                QKVParallelLinear, RowParallelLinear)
from sglang_security_b200.layers.quantization.awq import AWQConfig
from sglang_security_b200.layers.quantization.fp8 import Fp8Config
from sglang_security_b200.layers.quantization.gptq import GPTQConfig

_FIXTURE = torch.load(os.path.join(os.path.dirname(__file__), "golden", "weight_loaders.pt"),
                      weights_only=False)
GOLD = _FIXTURE["cases"]


def _quant_config(quant):
    if quant is None:
        return None
    if quant[0] == "awq":
        return AWQConfig(4, quant[1], True)
    if quant[0] == "gptq":
        return GPTQConfig(4, quant[1], False)
    if quant[0] == "fp8_tensor":
        return Fp8Config(is_checkpoint_fp8_serialized=True, activation_scheme="static")
    if quant[0] == "fp8_block":
        return Fp8Config(is_checkpoint_fp8_serialized=True, activation_scheme="dynamic",
                         weight_block_size=list(quant[1]))
    raise KeyError(quant)


def _layer(case, rank):
    a, tp = case["args"], case["tp"]
    kw = dict(bias=False, params_dtype=getattr(torch, case["dtype"].split(".")[-1]),
              quant_config=_quant_config(case["quant"]), tp_rank=rank, tp_size=tp)
    if case["kind"] == "merged":
        return MergedColumnParallelLinear(a["hidden"], list(a["outs"]), **kw)
    if case["kind"] == "qkv":
        return QKVParallelLinear(a["hidden"], a["d"], a["heads"], a["kv_heads"], **kw)
    if case["kind"] == "row":
        return RowParallelLinear(a["in_size"], a["out_size"], **kw)
    return ColumnParallelLinear(a["in_size"], a["out_size"], **kw)


def _bits(t):
    return t.view(torch.uint8) if t.dtype == torch.float8_e4m3fn else t


def _shard_id(text):
    return None if text == "None" else (int(text) if text.isdigit() else text)


@pytest.mark.parametrize("case", GOLD, ids=[f"{c['name']}-tp{c['tp']}" for c in GOLD])
def test_loaders_equal_reference(case):
    ckpt = {tuple(k.split("|")): v for k, v in case["checkpoint"].items()}
    order = list(dict.fromkeys(sid for sid, _ in ckpt))            # projection order of the checkpoint
    for rank, gold in enumerate(case["ranks"]):
        layer = _layer(case, rank)
        for (sid, pname), t in ckpt.items():
            p = getattr(layer, pname)
            if _shard_id(sid) is None:
                p.weight_loader(p, t.clone())
            else:
                p.weight_loader(p, t.clone(), _shard_id(sid))
        for pname, want in gold["expected"].items():
            got = getattr(layer, pname).data
            assert got.shape == want.shape and got.dtype == want.dtype, (pname, got.shape, want.shape)
            assert torch.equal(_bits(got), _bits(want)), (case["name"], rank, pname)
        # the same projections fused on disk (no shard id)
        if gold["fused_params"]:
            fused = _layer(case, rank)
            for pname in gold["fused_params"]:
                p = getattr(fused, pname)
                dim = getattr(p, "output_dim", 0)
                p.weight_loader(p, torch.cat([ckpt[(sid, pname)] for sid in order], dim=dim))
                assert torch.equal(_bits(p.data), _bits(gold["expected"][pname])), (case["name"], rank, pname, "fused")


def test_fixture_covers_the_loader_matrix():
    names = {c["name"] for c in GOLD}
    assert {"merged_plain", "merged_awq", "merged_fp8_tensor", "merged_fp8_block", "qkv_plain", "qkv_awq",
            "qkv_fp8_tensor", "qkv_fp8_block", "row_plain", "row_awq", "row_gptq", "row_fp8_tensor",
            "column_plain"} <= names
    assert any(c["name"] == "qkv_plain" and c["tp"] == 8 for c in GOLD)      # kv heads replicated 4x


def test_fp8_per_shard_scales_to_channelwise_equal_reference():
    """Fp8LinearMethod.process_weights_after_loading on an fp8-serialized per-tensor checkpoint: the
    per-shard scales become one scale per output channel exactly as the reference's
    convert_to_channelwise does (quantization/utils.py:68-88), the static input scale becomes its max."""
    from sglang_security_b200.layers.quantization.fp8_utils import convert_to_channelwise

    assert len(_FIXTURE["channelwise"]) == 4
    for c in _FIXTURE["channelwise"]:
        got = convert_to_channelwise(c["weight_scale"], c["logical_widths"])
        assert got.dtype == c["expected"].dtype and torch.equal(got, c["expected"])
    # through the layer: QKV at tp 2 (heads 8 / kv 2 / d 4 -> widths 16, 4, 4)
    cfg = Fp8Config(is_checkpoint_fp8_serialized=True, activation_scheme="static")
    layer = QKVParallelLinear(32, 4, 8, 2, bias=False, params_dtype=torch.bfloat16, quant_config=cfg,
                              tp_rank=0, tp_size=2)
    for sid, val in (("q", 0.5), ("k", 0.25), ("v", 2.0)):
        layer.weight_scale.weight_loader(layer.weight_scale, torch.tensor(val), sid)
        layer.input_scale.weight_loader(layer.input_scale, torch.tensor(val * 3), sid)
    layer.quant_method.process_weights_after_loading(layer)
    want = torch.tensor([0.5] * 16 + [0.25] * 4 + [2.0] * 4).reshape(24, 1)
    assert torch.equal(layer.weight_scale.data, want)
    assert float(layer.input_scale) == 6.0 and tuple(layer.weight.shape) == (32, 24)
