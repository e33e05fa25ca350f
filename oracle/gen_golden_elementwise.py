"""Generate tests/golden/elementwise_*.pt by running the UNMODIFIED reference RMSNorm / SiluAndMul
/ RotaryEmbedding classes (forward_native) on CPU.  Run in the authoring container only:
    python -m oracle.gen_golden_elementwise
TEST INFRASTRUCTURE ONLY."""
from __future__ import annotations

import os

import torch

from . import ref_loader

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests",
                          "golden")


def cases():
    RMSNorm, SiluAndMul, RotaryEmbedding = ref_loader.load_elementwise_layers()
    g = torch.Generator().manual_seed(7)
    out = {}
    # RMSNorm with and without residual (Qwen2.5-0.5B hidden 896, a 128-multiple)
    for name, (m, k, dt) in {"rms_896_bf16": (5, 896, torch.bfloat16),
                             "rms_1024_f16": (3, 1024, torch.float16)}.items():
        norm = RMSNorm(k, eps=1e-6)
        norm.weight.data = (1 + 0.1 * torch.randn(k, generator=g)).to(dt)
        x = torch.randn(m, k, generator=g).to(dt)
        r = torch.randn(m, k, generator=g).to(dt)
        y = norm.forward_native(x.clone())
        y2, r2 = norm.forward_native(x.clone(), r.clone())
        out[name] = dict(x=x, residual=r, weight=norm.weight.data.clone(), eps=1e-6, y=y,
                         y_res=y2, residual_out=r2)
    # SiluAndMul
    x = (torch.randn(4, 2 * 256, generator=g) * 2).to(torch.bfloat16)
    out["silu_256_bf16"] = dict(x=x, y=SiluAndMul().forward_native(x))
    # RoPE: neox (Llama/Qwen) full rotation, neox partial rotation, gpt-j style
    for name, (hq, hkv, d, rot, neox) in {"rope_neox_d64": (14, 2, 64, 64, True),
                                         "rope_neox_partial_d128": (4, 2, 128, 64, True),
                                         "rope_gptj_d128": (4, 4, 128, 128, False)}.items():
        rope = RotaryEmbedding(d, rot, 512, 10000, neox, torch.bfloat16)
        # CPU reference keeps the cache in the model dtype (rotary_embedding.py:79-81); the fp32
        # cache used on CUDA is stored too
        pos = torch.randint(0, 512, (6,), generator=g)
        q = torch.randn(6, hq * d, generator=g).to(torch.bfloat16)
        k = torch.randn(6, hkv * d, generator=g).to(torch.bfloat16)
        qo, ko = rope.forward_native(pos, q.clone(), k.clone())
        out[name] = dict(positions=pos, q=q, k=k, q_out=qo, k_out=ko,
                         cos_sin_cache=rope.cos_sin_cache.clone(), head_size=d, rotary_dim=rot,
                         is_neox=neox, base=10000, max_pos=512)
    # Llama-3.1 rope_scaling (config.json of Llama-3.1-8B/70B): cos/sin table of the reference's
    # get_rope(..., rope_scaling={'rope_type': 'llama3', ...}) -> Llama3RotaryEmbedding
    rope_mod = ref_loader.load_rope_module()
    scaling = {"rope_type": "llama3", "factor": 8.0, "low_freq_factor": 1.0,
               "high_freq_factor": 4.0, "original_max_position_embeddings": 8192}
    rope_mod._ROPE_DICT.clear()
    r3 = rope_mod.get_rope(128, 128, 2048, 500000.0, True, dict(scaling), torch.bfloat16)
    assert type(r3).__name__ == "Llama3RotaryEmbedding", type(r3)
    out["rope_llama3_d128"] = dict(cos_sin_cache=r3.cos_sin_cache.clone(), head_size=128,
                                   rotary_dim=128, max_pos=2048, base=500000.0, rope_scaling=scaling)
    return out


def main():
    os.makedirs(GOLDEN_DIR, exist_ok=True)
    for name, payload in cases().items():
        path = os.path.join(GOLDEN_DIR, f"elementwise_{name}.pt")
        torch.save(payload, path)
        print("wrote", path)


if __name__ == "__main__":
    main()
