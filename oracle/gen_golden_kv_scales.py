"""Generates tests/golden/kv_cache_scales.pt by running the reference's OWN host code for the FP8
KV-cache scaling factors, extracted unmodified (ast) from the read-only reference tree:
  kv_cache_scales_loader, QuantParamSchema, KVCacheQuantSchema, maybe_remap_kv_scale_name
      python/sglang/srt/model_loader/weight_utils.py:670-857
  BaseKVCacheMethod.create_weights / process_weights_after_loading
      python/sglang/srt/layers/quantization/kv_cache.py:17-87
on the reference's own fixtures test/srt/kv_cache_scales_{llama3_8b,llama3_1_8b,qwen2_1_5b}.json
(the inputs of test/srt/test_fp8_kvcache.py) and on malformed variants of them (wrong TP size, a
missing layer, wrong dtype, wrong model type, broken JSON, missing file), which the reference answers
with "no scales" (-> 1.0).  The JSON documents are stored in the fixture next to the answers, so the
test needs no access to /root/reference.
Run here (needs /root/reference and pydantic).  TEST INFRASTRUCTURE ONLY."""
import ast
import copy
import json
import logging
import os
import sys
import tempfile
from typing import Dict, Iterable, Optional, Tuple

import torch

REF = "/root/reference"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def extract(path, names, ns):
    """exec the named top-level classes / functions of `path`, in file order, inside `ns`."""
    nodes = [n for n in ast.parse(open(path).read()).body
             if isinstance(n, (ast.ClassDef, ast.FunctionDef)) and n.name in names]
    assert sorted(n.name for n in nodes) == sorted(names), [n.name for n in nodes]
    exec(compile(ast.Module(nodes, []), path, "exec"), ns)
    return ns


def load_reference():
    from pydantic import BaseModel, ConfigDict, ValidationInfo, model_validator

    warnings = []
    wu = {"json": json, "logger": logging.getLogger("ref"), "Optional": Optional, "Dict": Dict,
          "Iterable": Iterable, "Tuple": Tuple, "BaseModel": BaseModel, "ConfigDict": ConfigDict,
          "ValidationInfo": ValidationInfo, "model_validator": model_validator,
          "print_warning_once": warnings.append}
    extract(f"{REF}/python/sglang/srt/model_loader/weight_utils.py",
            ["KVCacheQuantSchema", "QuantParamSchema", "kv_cache_scales_loader",
             "maybe_remap_kv_scale_name"], wu)
    kv = {"torch": torch, "QuantizationConfig": object, "QuantizeMethodBase": object,
          "RadixAttention": object, "is_fp8_fnuz": lambda: False}
    extract(f"{REF}/python/sglang/srt/layers/quantization/kv_cache.py", ["BaseKVCacheMethod"], kv)
    return wu, kv["BaseKVCacheMethod"]


def loader_cases():
    """-> list of dict(name, text (file contents or None = missing file), tp_rank, tp_size, layers, model_type)"""
    docs = {n: json.load(open(f"{REF}/test/srt/kv_cache_scales_{n}.json"))
            for n in ("llama3_8b", "llama3_1_8b", "qwen2_1_5b")}
    cases = [dict(name=n, text=json.dumps(d), tp_rank=0, tp_size=1,
                  layers=len(d["kv_cache"]["scaling_factor"]["0"]), model_type=d["model_type"])
             for n, d in docs.items()]
    base = docs["llama3_8b"]
    two = copy.deepcopy(base)                       # a 2-rank document, ranks hold different scales
    two["kv_cache"]["scaling_factor"]["1"] = {k: round(v * 2, 4) for k, v in
                                              base["kv_cache"]["scaling_factor"]["0"].items()}
    cases.append(dict(name="tp2_rank1", text=json.dumps(two), tp_rank=1, tp_size=2, layers=32,
                      model_type="llama"))
    cases.append(dict(name="model_type_none", text=json.dumps(base), tp_rank=0, tp_size=1, layers=32,
                      model_type=None))
    # malformed: every one of these makes the reference return [] (scales default to 1.0)
    cases.append(dict(name="bad_tp_size", text=json.dumps(base), tp_rank=0, tp_size=2, layers=32,
                      model_type="llama"))
    cases.append(dict(name="bad_layer_count", text=json.dumps(base), tp_rank=0, tp_size=1, layers=40,
                      model_type="llama"))
    cases.append(dict(name="bad_model_type", text=json.dumps(base), tp_rank=0, tp_size=1, layers=32,
                      model_type="qwen2"))
    wrong_dtype = copy.deepcopy(base)
    wrong_dtype["kv_cache"]["dtype"] = "float8_e5m2"
    cases.append(dict(name="bad_dtype", text=json.dumps(wrong_dtype), tp_rank=0, tp_size=1, layers=32,
                      model_type="llama"))
    hole = copy.deepcopy(base)
    hole["kv_cache"]["scaling_factor"]["0"]["40"] = hole["kv_cache"]["scaling_factor"]["0"].pop("7")
    cases.append(dict(name="missing_layer", text=json.dumps(hole), tp_rank=0, tp_size=1, layers=32,
                      model_type="llama"))
    no_kv = {"model_type": "llama"}
    cases.append(dict(name="no_kv_cache_key", text=json.dumps(no_kv), tp_rank=0, tp_size=1, layers=32,
                      model_type="llama"))
    cases.append(dict(name="broken_json", text=json.dumps(base)[:-20], tp_rank=0, tp_size=1,
                      layers=32, model_type="llama"))
    cases.append(dict(name="missing_file", text=None, tp_rank=0, tp_size=1, layers=32,
                      model_type="llama"))
    return cases


REMAP_NAMES = ["model.layers.3.self_attn.kv_scale", "model.layers.3.self_attn.k_scale",
               "model.layers.3.self_attn.v_scale", "model.layers.3.self_attn.k_proj.k_scale",
               "model.layers.3.self_attn.v_proj.v_scale", "model.layers.9.self_attn.kv_scale",
               "model.layers.9.self_attn.k_scale", "model.layers.3.self_attn.qkv_proj.weight_scale",
               "model.layers.3.mlp.down_proj.weight"]
REMAP_PARAMS = ["model.layers.3.self_attn.attn.k_scale", "model.layers.3.self_attn.attn.v_scale",
                "model.layers.3.self_attn.qkv_proj.weight_scale", "model.layers.3.mlp.down_proj.weight"]
METHOD_SCALES = [(-1.0, -1.0), (0.5, 0.25), (0.0408, 0.0408), (0.75, -1.0), (2.5, 0.125)]


def main():
    logging.disable(logging.CRITICAL)               # the loader logs every malformed document
    wu, method_cls = load_reference()
    out = {"loader": [], "remap": [], "method": []}
    with tempfile.TemporaryDirectory() as tmp:
        for c in loader_cases():
            path = os.path.join(tmp, c["name"] + ".json")
            if c["text"] is not None:
                open(path, "w").write(c["text"])
            got = wu["kv_cache_scales_loader"](path, c["tp_rank"], c["tp_size"], c["layers"],
                                               c["model_type"])
            c["expected"] = [(int(k), float(v)) for k, v in got]
            out["loader"].append(c)
            print(c["name"], len(c["expected"]), c["expected"][:2])
    params = {n: None for n in REMAP_PARAMS}
    for name in REMAP_NAMES:
        out["remap"].append((name, wu["maybe_remap_kv_scale_name"](name, params)))
    out["remap_params"] = REMAP_PARAMS
    for ks, vs in METHOD_SCALES:
        layer = torch.nn.Module()
        m = method_cls(None)
        m.create_weights(layer)
        created = (float(layer.k_scale), float(layer.v_scale), layer.k_scale.dtype, layer.k_scale.requires_grad)
        layer.k_scale.data.fill_(ks)
        layer.v_scale.data.fill_(vs)
        with torch.no_grad():
            m.process_weights_after_loading(layer)
        out["method"].append(dict(loaded=(ks, vs), created=created,
                                  k_scale=float(layer.k_scale), v_scale=float(layer.v_scale),
                                  k_scale_float=layer.k_scale_float, v_scale_float=layer.v_scale_float))
        print("method", (ks, vs), "->", layer.k_scale_float, layer.v_scale_float)
    for r in out["remap"]:
        print("remap", r)
    path = os.path.join(ROOT, "tests", "golden", "kv_cache_scales.pt")
    torch.save(out, path)
    print(path, os.path.getsize(path))


if __name__ == "__main__":
    sys.exit(main())
