"""ORACLE (test infrastructure only — never imported by the product path).

CPU restatements of the reference's torch-native implementations of the ops either side of the
GEMMs / attention:
  rmsnorm_oracle        RMSNorm.forward_native, python/sglang/srt/layers/layernorm.py:88-108
  silu_and_mul_oracle   SiluAndMul.forward_native, python/sglang/srt/layers/activation.py:43-45
  rope_oracle           RotaryEmbedding.forward_native + _apply_rotary_emb,
                        python/sglang/srt/layers/rotary_embedding.py:30-57,119-147
  cos_sin_cache_oracle  RotaryEmbedding._compute_cos_sin_cache, rotary_embedding.py:93-117
Pinned by tests/golden/elementwise_*.pt = outputs of the UNMODIFIED reference classes loaded by
file path (oracle/gen_golden_elementwise.py)."""
import torch
import torch.nn.functional as F


def rmsnorm_oracle(x, weight, eps, residual=None):
    orig_dtype = x.dtype
    x = x.to(torch.float32)
    if residual is not None:
        x = x + residual.to(torch.float32)
        residual = x.to(orig_dtype)
    variance = x.pow(2).mean(dim=-1, keepdim=True)
    x = x * torch.rsqrt(variance + eps)
    x = (x * weight).to(orig_dtype)
    return x if residual is None else (x, residual)


def silu_and_mul_oracle(x):
    d = x.shape[-1] // 2
    return F.silu(x[..., :d]) * x[..., d:]


def cos_sin_cache_oracle(rotary_dim, max_pos, base):
    inv_freq = 1.0 / (base ** (torch.arange(0, rotary_dim, 2, dtype=torch.float) / rotary_dim))
    t = torch.arange(max_pos, dtype=torch.float)
    freqs = torch.einsum("i,j -> ij", t, inv_freq)
    return torch.cat((freqs.cos(), freqs.sin()), dim=-1)


def _apply_rotary_emb(x, cos, sin, is_neox_style):
    cos = cos.unsqueeze(-2).to(x.dtype)
    sin = sin.unsqueeze(-2).to(x.dtype)
    if is_neox_style:
        x1, x2 = torch.chunk(x, 2, dim=-1)
    else:
        x1 = x[..., ::2]
        x2 = x[..., 1::2]
    o1 = x1 * cos - x2 * sin
    o2 = x2 * cos + x1 * sin
    if is_neox_style:
        return torch.cat((o1, o2), dim=-1)
    return torch.stack((o1, o2), dim=-1).flatten(-2)


def rope_oracle(positions, query, key, cos_sin_cache, head_size, rotary_dim, is_neox_style):
    positions = positions.flatten()
    num_tokens = positions.shape[0]
    cos_sin = cos_sin_cache.index_select(0, positions)
    cos, sin = cos_sin.chunk(2, dim=-1)
    out = []
    for t in (query, key):
        shape = t.shape
        t = t.view(num_tokens, -1, head_size)
        rot = _apply_rotary_emb(t[..., :rotary_dim], cos, sin, is_neox_style)
        out.append(torch.cat((rot, t[..., rotary_dim:]), dim=-1).reshape(shape))
    return out[0], out[1]
