"""Load pieces of the UNMODIFIED reference by file path (authoring container only).

TEST INFRASTRUCTURE ONLY.  `/root/reference` does not exist on the GPU box, so nothing that runs
there may call this module; it is used by oracle/gen_golden.py to produce tests/golden/*.pt and by
the `not gpu` tests that re-check the restatement against the live reference when it is present.

`import sglang` itself fails in this image (pins torch 2.6 / triton 3.2 / transformers 4.52), so
single files are executed with the minimum of stub modules.
"""
from __future__ import annotations

import enum
import importlib.util
import os
import sys
import types

REF_ROOT = os.environ.get("B200_REFERENCE_ROOT", "/root/reference")
SRT = os.path.join(REF_ROOT, "python", "sglang", "srt")


def available() -> bool:
    return os.path.isfile(os.path.join(SRT, "layers", "attention", "torch_native_backend.py"))


def _stub(name: str):
    m = types.ModuleType(name)
    m.__path__ = []
    sys.modules[name] = m
    return m


def _exec(mod_name: str, path: str):
    spec = importlib.util.spec_from_file_location(mod_name, path)
    m = importlib.util.module_from_spec(spec)
    sys.modules[mod_name] = m
    spec.loader.exec_module(m)
    return m


_cached = None


def load_torch_native_backend():
    """Returns the reference module object of layers/attention/torch_native_backend.py."""
    global _cached
    if _cached is not None:
        return _cached
    if not available():
        raise RuntimeError(f"reference tree not found under {REF_ROOT}")
    saved = {k: v for k, v in sys.modules.items() if k == "sglang" or k.startswith("sglang.")}
    try:
        for n in ["sglang", "sglang.srt", "sglang.srt.layers", "sglang.srt.layers.attention",
                  "sglang.srt.model_executor"]:
            _stub(n)
        ra = _stub("sglang.srt.layers.radix_attention")

        class AttentionType(enum.Enum):  # values as in layers/radix_attention.py:26-35
            DECODER = "decoder"
            ENCODER_ONLY = "encoder_only"

        ra.AttentionType = AttentionType
        fb = _stub("sglang.srt.model_executor.forward_batch_info")
        fb.ForwardBatch = object
        fb.ForwardMode = object
        _exec("sglang.srt.layers.attention.base_attn_backend",
              os.path.join(SRT, "layers", "attention", "base_attn_backend.py"))
        mod = _exec("sglang.srt.layers.attention.torch_native_backend",
                    os.path.join(SRT, "layers", "attention", "torch_native_backend.py"))
        mod.AttentionType = AttentionType
    finally:
        for k in [k for k in sys.modules if k == "sglang" or k.startswith("sglang.")]:
            del sys.modules[k]
        sys.modules.update(saved)
    _cached = mod
    return mod


def make_reference_backend():
    mod = load_torch_native_backend()
    return mod.TorchNativeAttnBackend(types.SimpleNamespace(device="cpu")), mod.AttentionType


def load_test_module(rel_path: str, name: str):
    """Execute a reference test file (e.g. sgl-kernel/tests/test_awq_dequant.py) far enough to get
    its torch reference functions.  Missing third-party imports are stubbed."""
    path = os.path.join(REF_ROOT, rel_path)
    src = open(path).read()
    mod = types.ModuleType(name)
    mod.__file__ = path
    return mod, src


_elementwise = None


def load_elementwise_layers():
    """Returns (RMSNorm, SiluAndMul, RotaryEmbedding) classes of the UNMODIFIED reference
    (layers/layernorm.py, layers/activation.py, layers/rotary_embedding.py + custom_op.py), loaded
    by file path.  `is_cuda()` is stubbed True (otherwise layernorm.py swaps its own RMSNorm for
    vllm's at the bottom of the file) with an empty `sgl_kernel` stub for the CUDA-op imports; the
    callers invoke `forward_native` explicitly — the torch-native implementation the north star
    names.  With is_cuda True the rotary cos/sin cache stays fp32 (rotary_embedding.py:79-81)."""
    global _elementwise
    if _elementwise is not None:
        return _elementwise
    if not available():
        raise RuntimeError(f"reference tree not found under {REF_ROOT}")
    saved = {k: v for k, v in sys.modules.items() if k == "sglang" or k.startswith("sglang.")}
    try:
        for n in ["sglang", "sglang.srt", "sglang.srt.layers", "sglang.srt.layers.quantization"]:
            _stub(n)
        utils = _stub("sglang.srt.utils")
        utils.is_cuda = lambda: True
        utils.is_hip = lambda: False
        saved_sgl = sys.modules.get("sgl_kernel")
        sk = types.ModuleType("sgl_kernel")
        for fn in ("fused_add_rmsnorm", "gemma_fused_add_rmsnorm", "gemma_rmsnorm", "rmsnorm",
                   "gelu_and_mul", "gelu_tanh_and_mul", "silu_and_mul",
                   "apply_rope_with_cos_sin_cache_inplace"):
            setattr(sk, fn, None)
        sys.modules["sgl_kernel"] = sk
        utils.set_weight_attrs = lambda *a, **k: None
        dist = _stub("sglang.srt.distributed")
        dist.divide = lambda a, b: a // b
        dist.get_tensor_model_parallel_rank = lambda: 0
        dist.get_tensor_model_parallel_world_size = lambda: 1
        bc = _stub("sglang.srt.layers.quantization.base_config")
        bc.QuantizationConfig = object
        _exec("sglang.srt.custom_op", os.path.join(SRT, "custom_op.py"))
        ln = _exec("sglang.srt.layers.layernorm", os.path.join(SRT, "layers", "layernorm.py"))
        act = _exec("sglang.srt.layers.activation", os.path.join(SRT, "layers", "activation.py"))
        rope = _exec("sglang.srt.layers.rotary_embedding",
                     os.path.join(SRT, "layers", "rotary_embedding.py"))
    finally:
        for k in [k for k in sys.modules if k == "sglang" or k.startswith("sglang.")]:
            del sys.modules[k]
        sys.modules.update(saved)
        if saved_sgl is not None:
            sys.modules["sgl_kernel"] = saved_sgl
        else:
            sys.modules.pop("sgl_kernel", None)
    _elementwise = (ln.RMSNorm, act.SiluAndMul, rope.RotaryEmbedding)
    global _rope_module
    _rope_module = rope
    return _elementwise


_rope_module = None


def load_rope_module():
    """The UNMODIFIED reference layers/rotary_embedding.py module (get_rope, Llama3RotaryEmbedding)."""
    load_elementwise_layers()
    return _rope_module
