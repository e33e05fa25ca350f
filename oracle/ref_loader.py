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
