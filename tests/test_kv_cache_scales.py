"""CPU: host side of the FP8 KV-cache scaling factors (kv_cache_scales_loader,
maybe_remap_kv_scale_name, BaseKVCacheMethod, load_kv_cache_scales) against the answers of the
reference's OWN functions on its own fixtures test/srt/kv_cache_scales_*.json and on malformed
variants (tests/golden/kv_cache_scales.pt, oracle/gen_golden_kv_scales.py)."""
import os

import pytest
import torch

from sglang_security_b200.layers.quantization.kv_cache import BaseKVCacheMethod, Fp8KVCacheMethod
from sglang_security_b200.layers.radix_attention import RadixAttention
from sglang_security_b200.model_loader.weight_utils import (kv_cache_scales_loader, load_kv_cache_scales,
                                                            maybe_remap_kv_scale_name)

GOLD = torch.load(os.path.join(os.path.dirname(__file__), "golden", "kv_cache_scales.pt"),
                  weights_only=False)


def _write(tmp_path, case):
    path = tmp_path / (case["name"] + ".json")
    if case["text"] is not None:
        path.write_text(case["text"])
    return str(path)


@pytest.mark.parametrize("case", GOLD["loader"], ids=[c["name"] for c in GOLD["loader"]])
def test_kv_cache_scales_loader_equals_reference(case, tmp_path):
    got = kv_cache_scales_loader(_write(tmp_path, case), case["tp_rank"], case["tp_size"],
                                 case["layers"], case["model_type"])
    assert [(int(k), float(v)) for k, v in got] == case["expected"]


def test_loader_cases_cover_good_and_malformed_documents():
    names = {c["name"]: len(c["expected"]) for c in GOLD["loader"]}
    assert names["llama3_8b"] == 32 and names["qwen2_1_5b"] == 28 and names["tp2_rank1"] == 32
    for bad in ("bad_tp_size", "bad_layer_count", "bad_model_type", "bad_dtype", "missing_layer",
                "no_kv_cache_key", "broken_json", "missing_file"):
        assert names[bad] == 0


def test_remap_kv_scale_name_equals_reference():
    params = {n: None for n in GOLD["remap_params"]}
    assert len(GOLD["remap"]) == 9
    for name, expected in GOLD["remap"]:
        assert maybe_remap_kv_scale_name(name, params) == expected, name


@pytest.mark.parametrize("g", GOLD["method"], ids=[str(g["loaded"]) for g in GOLD["method"]])
def test_kv_cache_method_equals_reference(g):
    class Cfg:
        def get_quant_method(self, layer, prefix=""):
            return Fp8KVCacheMethod(self) if isinstance(layer, RadixAttention) else None

    layer = RadixAttention(8, 128, 128 ** -0.5, 2, 0, quant_config=Cfg(), prefix="model.layers.0.self_attn.attn")
    assert isinstance(layer.quant_method, BaseKVCacheMethod)
    assert (float(layer.k_scale), float(layer.v_scale), layer.k_scale.dtype,
            layer.k_scale.requires_grad) == g["created"]
    assert sorted(n for n, _ in layer.named_parameters()) == ["k_scale", "v_scale"]   # loadable by name
    layer.k_scale.data.fill_(g["loaded"][0])
    layer.v_scale.data.fill_(g["loaded"][1])
    layer.quant_method.process_weights_after_loading(layer)
    assert float(layer.k_scale) == g["k_scale"] and float(layer.v_scale) == g["v_scale"]
    assert layer.k_scale_float == g["k_scale_float"] and layer.v_scale_float == g["v_scale_float"]
    with pytest.raises(RuntimeError):
        layer.quant_method.apply(layer)


def test_load_kv_cache_scales_sets_every_layer(tmp_path):
    case = next(c for c in GOLD["loader"] if c["name"] == "llama3_8b")
    layers = [RadixAttention(8, 128, 128 ** -0.5, 2, i) for i in range(32)]
    assert load_kv_cache_scales(layers, _write(tmp_path, case), 0, 1, "llama") == 32
    for i, s in case["expected"]:
        a = layers[i]
        assert a.k_scale == s and a.v_scale == s and a.k_scale_float == s and a.v_scale_float == s
    # a malformed document leaves the defaults (scale-free pool writes, 1.0 in attention)
    fresh = [RadixAttention(8, 128, 128 ** -0.5, 2, i) for i in range(32)]
    bad = next(c for c in GOLD["loader"] if c["name"] == "bad_dtype")
    assert load_kv_cache_scales(fresh, _write(tmp_path, bad), 0, 1, "llama") == 0
    assert all(a.k_scale is None and a.k_scale_float is None for a in fresh)
    # layers whose scales are parameters (KV-cache quant method) are filled in place
    class Cfg:
        def get_quant_method(self, layer, prefix=""):
            return Fp8KVCacheMethod(self)

    par = [RadixAttention(8, 128, 128 ** -0.5, 2, i, quant_config=Cfg()) for i in range(32)]
    assert load_kv_cache_scales(par, _write(tmp_path, case), 0, 1, None) == 32
    assert isinstance(par[5].k_scale, torch.nn.Parameter)
    assert float(par[5].k_scale) == pytest.approx(case["expected"][5][1]) and par[5].v_scale_float == case["expected"][5][1]
