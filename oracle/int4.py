"""CPU restatements of the INT4 weight-only formats (AWQ, GPTQ).  TEST INFRASTRUCTURE ONLY.

AWQ follows awq_dequantize_torch (sgl-kernel/tests/test_awq_dequant.py:9-57) and
AWQLinearMethod.apply (python/sglang/srt/layers/quantization/awq.py:197-215).

GPTQ: the arithmetic lives in vllm 0.8.4 (`vllm...gptq.GPTQLinearMethod`, `ops.gptq_gemm`), which is
NOT under /root/reference => PARITY UNPINNED.  Restated from the published AutoGPTQ v1 checkpoint
format (format knowledge, not reference source): qweight [K/8, N] int32 packs 8 consecutive k per
word (nibble k%8), qzeros [K/g, N/8] packs 8 consecutive n per word and stores zero-1,
w[k,n] = (q - (z + 1)) * scales[g_idx[k], n].
"""
from __future__ import annotations

import torch

AWQ_ORDER = [0, 4, 1, 5, 2, 6, 3, 7]  # output column j of a word sits at nibble AWQ_ORDER[j]


def unpack_awq_int4(packed: torch.Tensor) -> torch.Tensor:
    """[R, C] int32 -> [R, 8C] integers 0..15 in output-column order (bit-exact target)."""
    shifts = torch.tensor([4 * o for o in AWQ_ORDER], dtype=torch.int32, device=packed.device)
    nib = (packed.unsqueeze(-1) >> shifts) & 0xF
    return nib.reshape(packed.shape[0], -1)


def awq_dequantize_oracle(qweight, scales, qzeros, group_size=None):
    """-> [K, N] in scales.dtype:  (w - z) * s  with the product rounded once to 16 bit."""
    k = qweight.shape[0]
    if group_size is None or group_size == -1:
        group_size = k // scales.shape[0]
    w = unpack_awq_int4(qweight)
    z = unpack_awq_int4(qzeros).repeat_interleave(group_size, dim=0)
    s = scales.repeat_interleave(group_size, dim=0)
    return ((w - z).to(torch.float32) * s.to(torch.float32)).to(scales.dtype)


def unpack_gptq_qweight(qweight: torch.Tensor) -> torch.Tensor:
    """[K/8, N] int32 -> [K, N] integers 0..15 (nibble i of word r is row 8r + i)."""
    shifts = torch.arange(0, 32, 4, dtype=torch.int32)
    nib = (qweight.unsqueeze(1) >> shifts.view(1, 8, 1)) & 0xF
    return nib.reshape(-1, qweight.shape[1])


def unpack_gptq_qzeros(qzeros: torch.Tensor) -> torch.Tensor:
    """[G, N/8] int32 -> [G, N] effective zero points (stored value + 1)."""
    shifts = torch.arange(0, 32, 4, dtype=torch.int32)
    nib = (qzeros.unsqueeze(-1) >> shifts) & 0xF
    return nib.reshape(qzeros.shape[0], -1) + 1


def gptq_dequantize_oracle(qweight, scales, qzeros, g_idx=None, group_size=128):
    w = unpack_gptq_qweight(qweight)
    z = unpack_gptq_qzeros(qzeros)
    k = w.shape[0]
    if g_idx is None:
        g_idx = torch.arange(k) // group_size
    g_idx = g_idx.long()
    return ((w - z[g_idx]).to(torch.float32) * scales[g_idx].to(torch.float32)).to(scales.dtype)


def w4a16_linear_oracle(x, w_dequant, bias=None):
    """matmul of the 16-bit dequantised weight with fp32 accumulation, output in x.dtype."""
    out = (x.reshape(-1, x.shape[-1]).float() @ w_dequant.float()).to(x.dtype)
    if bias is not None:
        out = out + bias
    return out.reshape(*x.shape[:-1], w_dequant.shape[1])


def gemv_stream_layout_oracle(qweight: torch.Tensor) -> torch.Tensor:
    """B200-internal stream layout of the decode GEMV (csrc/w4a16_gemv2.cu; no reference twin —
    the reference never repacks AWQ weights): [K, N/8] packed words -> flat int32 words ordered
    [n / 16][k / 64][lane = 4 g + t][j], word j = the 8 weights of column 16 nt + g + 8 (j & 1) at
    k = 64 kb + 32 (j >> 1) + 8 t + {0, 2, 4, 6, 1, 3, 5, 7} (nibble order).  Pure index shuffling,
    so the check is bit-exact."""
    ints = unpack_awq_int4(qweight).to(torch.int64)            # [K, N]
    k, n = ints.shape
    assert k % 64 == 0 and n % 16 == 0
    nt, kb, g, t, j, i = torch.meshgrid(torch.arange(n // 16), torch.arange(k // 64), torch.arange(8),
                                        torch.arange(4), torch.arange(4), torch.arange(8),
                                        indexing="ij")
    col = nt * 16 + g + 8 * (j & 1)
    kk = kb * 64 + (j >> 1) * 32 + 8 * t + torch.where(i < 4, 2 * i, 2 * (i - 4) + 1)
    words = (ints[kk, col] << (4 * i)).sum(-1)                 # [nt, kb, g, t, j]
    words = torch.where(words >= 2 ** 31, words - 2 ** 32, words)
    return words.reshape(-1).to(torch.int32)
