"""RotaryEmbedding with the reference's interface (python/sglang/srt/layers/rotary_embedding.py:
59-176): same constructor and cos_sin_cache construction (:93-117, fp32 cache on CUDA :79-81),
`forward(positions, query, key)` rotates query/key IN PLACE like the reference's CUDA path
(apply_rope_with_cos_sin_cache_inplace, :156-164) and returns them.

`forward_with_kv_write` fuses MHATokenToKVPool.set_kv_buffer (memory_pool.py:378-412) of the
rotated key and the value into the same kernel (b200_rope_kv_write); a model that uses it calls the
attention layer with save_kv_cache=False."""
from __future__ import annotations

import math
from typing import Optional, Tuple

import torch
import torch.nn as nn

from .. import _lib


class RotaryEmbedding(nn.Module):
    def __init__(self, head_size: int, rotary_dim: int, max_position_embeddings: int, base: float,
                 is_neox_style: bool, dtype: torch.dtype, native_rounding: bool = False) -> None:
        super().__init__()
        self.head_size = head_size
        self.rotary_dim = rotary_dim
        self.max_position_embeddings = max_position_embeddings
        self.base = base
        self.is_neox_style = is_neox_style
        self.dtype = dtype
        # True: bit-exact with forward_native's 16-bit arithmetic; False: fp32 math, one rounding
        # (the reference's CUDA kernel)
        self.native_rounding = native_rounding
        self.register_buffer("cos_sin_cache", self._compute_cos_sin_cache(), persistent=False)

    def _compute_inv_freq(self, base: float) -> torch.Tensor:
        return 1.0 / (base ** (torch.arange(0, self.rotary_dim, 2, dtype=torch.float)
                               / self.rotary_dim))

    def _compute_cos_sin_cache(self) -> torch.Tensor:
        inv_freq = self._compute_inv_freq(self.base)
        t = torch.arange(self.max_position_embeddings, dtype=torch.float)
        freqs = torch.einsum("i,j -> ij", t, inv_freq)
        return torch.cat((freqs.cos(), freqs.sin()), dim=-1)

    def _launch(self, positions, query, key, value, k_cache, v_cache, loc, fp8_scales=None,
                pool_dtype=None):
        if not query.is_cuda:
            raise _lib.B200Error("RotaryEmbedding: tensors must live on a CUDA device (no CPU path)")
        t = positions.numel()
        q2 = query.view(t, -1)
        k2 = key.view(t, -1)
        hq = q2.shape[1] // self.head_size
        hkv = k2.shape[1] // self.head_size
        if q2.stride(-1) != 1 or k2.stride(-1) != 1:
            raise _lib.B200Error("RotaryEmbedding: query/key rows must be contiguous")
        pos = positions.reshape(-1)
        if pos.dtype != torch.int64:
            pos = pos.to(torch.int64)
        cache = self.cos_sin_cache
        if cache.device != query.device:
            cache = cache.to(query.device)
            self.cos_sin_cache = cache
        v2 = value.view(t, -1) if value is not None else None
        if fp8_scales is not None:  # FP8 pool: rotate + quantising scatter in one launch
            ks, vs = fp8_scales
            rc = _lib.load().b200_rope_kv_write_fp8(
                pos.data_ptr(), q2.data_ptr(), q2.stride(0), k2.data_ptr(), k2.stride(0),
                v2.data_ptr(), v2.stride(0), cache.data_ptr(), _lib.dtype_code(cache.dtype),
                self.rotary_dim, 1 if self.is_neox_style else 0, 1 if self.native_rounding else 0,
                k_cache.data_ptr(), v_cache.data_ptr(), k_cache.stride(0), loc.data_ptr(),
                float(ks) if ks is not None else 1.0, float(vs) if vs is not None else 1.0,
                0 if ks is None else 1, t, hq, hkv, self.head_size,
                _lib.kv_dtype_code(query.dtype, pool_dtype), _lib.current_stream())
            _lib.check(rc, "b200_rope_kv_write_fp8")
            return
        rc = _lib.load().b200_rope_kv_write(
            pos.data_ptr(), q2.data_ptr(), q2.stride(0), k2.data_ptr(), k2.stride(0),
            _lib.ptr(v2), v2.stride(0) if v2 is not None else 0, cache.data_ptr(),
            _lib.dtype_code(cache.dtype), self.rotary_dim, 1 if self.is_neox_style else 0,
            1 if self.native_rounding else 0, _lib.ptr(k_cache), _lib.ptr(v_cache),
            k_cache.stride(0) if k_cache is not None else 0, _lib.ptr(loc), t, hq, hkv,
            self.head_size, _lib.dtype_code(query.dtype), _lib.current_stream())
        _lib.check(rc, "b200_rope_kv_write")

    def forward(self, positions: torch.Tensor, query: torch.Tensor, key: torch.Tensor,
                offsets: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, torch.Tensor]:
        if offsets is not None:
            positions = positions + offsets
        self._launch(positions, query, key, None, None, None, None)
        return query, key

    def forward_with_kv_write(self, positions, query, key, value, layer, forward_batch):
        """Rotate q/k in place and write (k, v) of every token to the KV pool rows
        forward_batch.out_cache_loc — the buffers are fetched through get_key_buffer /
        get_value_buffer per call so pool-side hooks keep firing."""
        pool = forward_batch.token_to_kv_pool
        k_cache = pool.get_key_buffer(layer.layer_id)
        v_cache = pool.get_value_buffer(layer.layer_id)
        loc = forward_batch.out_cache_loc
        if loc.dtype != torch.int64:
            loc = loc.to(torch.int64)
        if _lib.is_fp8_kv(k_cache.dtype):
            # FP8 pool: the quantising scatter of b200_kv_write_fp8 inside the RoPE kernel; scales as
            # in FlashInfer's set_kv_buffer call (flashinfer_backend.py:427), given together or not at all
            ks = getattr(layer, "k_scale_float", None)
            vs = getattr(layer, "v_scale_float", None)
            if (ks is None) != (vs is None):
                raise _lib.B200Error("forward_with_kv_write: k_scale and v_scale must be given together")
            if not k_cache.is_contiguous() or not v_cache.is_contiguous():
                raise _lib.B200Error("forward_with_kv_write: FP8 KV pool rows must be contiguous")
            self._launch(positions, query, key, value,
                         k_cache.view(k_cache.shape[0], -1).view(torch.uint8),
                         v_cache.view(v_cache.shape[0], -1).view(torch.uint8), loc, (ks, vs),
                         k_cache.dtype)
            return query, key
        if k_cache.dtype != key.dtype:
            raise _lib.B200Error("forward_with_kv_write: KV pool dtype differs from the activations")
        self._launch(positions, query, key, value, k_cache.view(k_cache.shape[0], -1),
                     v_cache.view(v_cache.shape[0], -1), loc)
        return query, key


class Llama3RotaryEmbedding(RotaryEmbedding):
    """Llama-3.1 frequency rescale (rotary_embedding.py:706-750): wavelengths shorter than
    orig_max/high_freq_factor keep their frequency, longer than orig_max/low_freq_factor are divided
    by `scaling_factor`, the band between is interpolated.  Only the cos/sin table changes; the
    kernels are the same."""

    def __init__(self, head_size: int, rotary_dim: int, max_position_embeddings: int, base: float,
                 is_neox_style: bool, dtype: torch.dtype, scaling_factor: float,
                 low_freq_factor: float, high_freq_factor: float, orig_max_position: int,
                 native_rounding: bool = False) -> None:
        self.scaling_factor = scaling_factor
        self.low_freq_factor = low_freq_factor
        self.high_freq_factor = high_freq_factor
        self.orig_max_position = orig_max_position
        super().__init__(head_size, rotary_dim, max_position_embeddings, base, is_neox_style, dtype,
                         native_rounding)

    def _compute_inv_freq(self, base: float) -> torch.Tensor:
        inv_freq = super()._compute_inv_freq(base)
        wave_len = 2 * math.pi / inv_freq
        lo_len = self.orig_max_position / self.low_freq_factor
        hi_len = self.orig_max_position / self.high_freq_factor
        smooth = 0
        if self.low_freq_factor != self.high_freq_factor:
            smooth = ((self.orig_max_position / wave_len - self.low_freq_factor)
                      / (self.high_freq_factor - self.low_freq_factor))
        scaled = inv_freq / self.scaling_factor
        mid = (1 - smooth) * scaled + smooth * inv_freq
        return torch.where(wave_len < hi_len, inv_freq, torch.where(wave_len > lo_len, scaled, mid))


def get_rope(head_size: int, rotary_dim: int, max_position: int, base: float,
             is_neox_style: bool = True, rope_scaling: Optional[dict] = None,
             dtype: torch.dtype = torch.bfloat16) -> RotaryEmbedding:
    """Subset of the reference factory (rotary_embedding.py:1083-1300 get_rope): unscaled RoPE and
    the `llama3` scaling the BASELINE models (Llama-3.1-8B / 70B) use.  Every other rope_scaling
    type raises instead of silently building unscaled tables."""
    if rope_scaling is None:
        return RotaryEmbedding(head_size, rotary_dim, max_position, base, is_neox_style, dtype)
    kind = rope_scaling.get("rope_type", rope_scaling.get("type"))
    if kind == "default":
        return RotaryEmbedding(head_size, rotary_dim, max_position, base, is_neox_style, dtype)
    if kind == "llama3":
        return Llama3RotaryEmbedding(head_size, rotary_dim, max_position, base, is_neox_style, dtype,
                                     rope_scaling["factor"], rope_scaling["low_freq_factor"],
                                     rope_scaling["high_freq_factor"],
                                     rope_scaling["original_max_position_embeddings"])
    raise NotImplementedError(f"rope_scaling type {kind!r} is not built (only None/'default'/'llama3')")
