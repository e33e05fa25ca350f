"""Generates tests/golden/triton_attn.pt by running the reference's OWN Triton attention kernels —
the production path of its default backend — through Triton's CPU interpreter (TRITON_INTERPRET=1):
  decode_attention_fwd  python/sglang/srt/layers/attention/triton_ops/decode_attention.py:682-735
                        (stage-1 split-KV kernels :50-175 / :246-407, stage-2 reduction :497-558)
  extend_attention_fwd  python/sglang/srt/layers/attention/triton_ops/extend_attention.py:306-438
The modules are imported UNMODIFIED by file path; their only sglang imports (`sglang.srt.utils`
is_cuda / is_hip, and prefill_attention.py for the extend module) are answered by a stub module and a
path-loaded module.  They are called the way TritonAttnBackend.forward_decode / forward_extend call
them (triton_backend.py:742-788, 790-835): new K/V rows written to the pool first, kv_indptr /
kv_indices over the whole sequence (decode) or the cached prefix (extend).
All scenario cases of oracle/scenarios.py are run in float16 (the interpreter is numpy-backed and has
no bfloat16), decode with per-request split counts as get_num_kv_splits produces them, plus one
decode and one extend case with logit capping.
Run here (needs /root/reference and triton).  TEST INFRASTRUCTURE ONLY."""
import importlib.util
import os
import sys
import types

os.environ["TRITON_INTERPRET"] = "1"   # before triton is imported

import torch

REF = "/root/reference/python/sglang/srt/layers/attention/triton_ops"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import scenarios  # noqa: E402
from oracle.pools import kv_indices_oracle  # noqa: E402

MAX_KV_SPLITS = 8


def load_reference_modules():
    for name in ("sglang", "sglang.srt", "sglang.srt.layers", "sglang.srt.layers.attention",
                 "sglang.srt.layers.attention.triton_ops"):
        sys.modules.setdefault(name, types.ModuleType(name))
    utils = types.ModuleType("sglang.srt.utils")
    utils.is_hip = lambda: False
    utils.is_cuda = lambda: False      # no device here: the kernels take their default tile sizes
    sys.modules["sglang.srt.utils"] = utils
    mods = {}
    for short in ("prefill_attention", "decode_attention", "extend_attention"):
        full = f"sglang.srt.layers.attention.triton_ops.{short}"
        spec = importlib.util.spec_from_file_location(full, os.path.join(REF, short + ".py"))
        mod = importlib.util.module_from_spec(spec)
        sys.modules[full] = mod
        spec.loader.exec_module(mod)
        mods[short] = mod
    return mods["decode_attention"], mods["extend_attention"]


def all_cases():
    cases = {}
    for n, kw in {**scenarios.DECODE_CASES, **scenarios.EXTEND_CASES}.items():
        cases[n] = dict(kw, dtype=torch.float16)
    cases["decode_qwen_g7_d64_cap"] = dict(scenarios.DECODE_CASES["decode_qwen_g7_d64"],
                                           dtype=torch.float16, logit_cap=3.0)
    cases["extend_llama_g4_d128_cap"] = dict(scenarios.EXTEND_CASES["extend_llama_g4_d128"],
                                             dtype=torch.float16, logit_cap=2.0)
    return cases


def run_case(dec, ext, kw):
    kw = dict(kw)
    logit_cap = kw.pop("logit_cap", 0.0)
    case = scenarios.make_attention_case(**kw)
    kc, vc = case.k_cache.clone(), case.v_cache.clone()
    kc[case.out_cache_loc] = case.k_new          # set_kv_buffer (pinned separately, bit-exact)
    vc[case.out_cache_loc] = case.v_new
    t = case.q.shape[0]
    q = case.q.view(t, case.hq, case.d)
    o = torch.zeros_like(q)
    scale = case.d ** -0.5
    extra = {}
    if case.decode:
        indptr, indices = kv_indices_oracle(case.req_to_token, case.req_pool_indices, case.seq_lens)
        logits = torch.zeros((case.bs, case.hq, MAX_KV_SPLITS, case.d), dtype=torch.float32)
        lse = torch.zeros((case.bs, case.hq, MAX_KV_SPLITS), dtype=torch.float32)
        # 1..MAX_KV_SPLITS splits, more for longer requests (get_num_kv_splits, triton_backend.py:40-73)
        splits = torch.clamp((case.seq_lens + 63) // 64, 1, MAX_KV_SPLITS).to(torch.int32)
        dec.decode_attention_fwd(q, kc, vc, o, indptr, indices, logits, lse, splits, MAX_KV_SPLITS,
                                 scale, logit_cap)
        extra["num_kv_splits"] = splits
    else:
        pre, extl = case.extend_prefix_lens, case.extend_seq_lens
        indptr, indices = kv_indices_oracle(case.req_to_token, case.req_pool_indices, pre)
        qo = torch.zeros(case.bs + 1, dtype=torch.int32)
        qo[1:] = torch.cumsum(extl, 0)
        ext.extend_attention_fwd(q, case.k_new.contiguous(), case.v_new.contiguous(), o, kc, vc, qo,
                                 indptr, indices, None, True, None, int(extl.max()), scale, logit_cap)
    return case, o.reshape(t, -1), logit_cap, extra


def main():
    dec, ext = load_reference_modules()
    out = {}
    for name, kw in all_cases().items():
        case, o, cap, extra = run_case(dec, ext, kw)
        assert torch.isfinite(o.float()).all(), name
        out[name] = {"kwargs": {k: (str(v) if isinstance(v, torch.dtype) else v) for k, v in kw.items()},
                     "input_digest": scenarios.digest(case), "logit_cap": cap, "o": o, **extra}
        print(name, tuple(o.shape), float(o.float().abs().max()))
    path = os.path.join(ROOT, "tests", "golden", "triton_attn.pt")
    torch.save({"cases": out, "source": "reference triton_ops decode/extend kernels, Triton CPU "
                "interpreter, torch " + torch.__version__}, path)
    print(path, os.path.getsize(path))


if __name__ == "__main__":
    sys.exit(main())
