"""FP8 kernels of the linear path, bound through the C-ABI.  Function names and argument meaning
follow the reference wrappers (python/sglang/srt/layers/quantization/fp8_kernel.py and
sgl-kernel/python/sgl_kernel/gemm.py) so call sites read the same."""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from ... import _lib

fp8_dtype = torch.float8_e4m3fn
fp8_max = torch.finfo(fp8_dtype).max
fp8_min = -fp8_max


def _check_2d_cuda(x: torch.Tensor, what: str):
    if not x.is_cuda:
        raise _lib.B200Error(f"{what}: tensor must live on a CUDA device (no CPU path)")
    if x.dim() != 2 or x.stride(-1) != 1:
        raise _lib.B200Error(f"{what}: expected a 2-D tensor with contiguous rows")


def sglang_per_token_quant_fp8(x: torch.Tensor, dtype: torch.dtype = fp8_dtype
                               ) -> Tuple[torch.Tensor, torch.Tensor]:
    """Dynamic per-token quantisation -> (x_q [M,K] e4m3, x_s [M,1] fp32)   (fp8_kernel.py:317-333)."""
    _check_2d_cuda(x, "sglang_per_token_quant_fp8")
    m, k = x.shape
    x_q = torch.empty((m, k), device=x.device, dtype=dtype)
    x_s = torch.empty((m, 1), device=x.device, dtype=torch.float32)
    rc = _lib.load().b200_fp8_quant_per_token(x.data_ptr(), x.stride(0), x_q.data_ptr(), k,
                                               x_s.data_ptr(), None, m, k, 0, 0.0,
                                               _lib.dtype_code(x.dtype), _lib.current_stream())
    _lib.check(rc, "b200_fp8_quant_per_token")
    return x_q, x_s


def static_quant_fp8(x: torch.Tensor, x_s: torch.Tensor, repeat_scale: bool = False
                     ) -> Tuple[torch.Tensor, torch.Tensor]:
    """Static per-tensor quantisation with the given scalar scale (fp8_kernel.py:379-432)."""
    _check_2d_cuda(x, "static_quant_fp8")
    assert x_s.numel() == 1, "only static per-tensor quantization is supported"
    m, k = x.shape
    x_q = torch.empty((m, k), device=x.device, dtype=fp8_dtype)
    s = x_s.reshape(1).to(torch.float32)
    rc = _lib.load().b200_fp8_quant_per_token(x.data_ptr(), x.stride(0), x_q.data_ptr(), k, None,
                                               s.data_ptr(), m, k, 0, 0.0, _lib.dtype_code(x.dtype),
                                               _lib.current_stream())
    _lib.check(rc, "b200_fp8_quant_per_token(static)")
    if repeat_scale:
        return x_q, s.reshape(1, 1).expand(m, 1).contiguous()
    return x_q, x_s


def per_token_group_quant_fp8(x: torch.Tensor, group_size: int, eps: float = 1e-10,
                              column_major_scales: bool = False, scale_tma_aligned: bool = False
                              ) -> Tuple[torch.Tensor, torch.Tensor]:
    """Per-(token, group) quantisation -> (x_q, x_s [M, K/group]) (fp8_kernel.py:190-274).
    group_size == K gives the per-channel weight quantisation used at load time (fp8.py:340-343)."""
    _check_2d_cuda(x, "per_token_group_quant_fp8")
    m, k = x.shape
    assert k % group_size == 0
    if group_size != 128:
        if group_size == k:  # one group per row, division semantics (load-time weight quant)
            x_q = torch.empty((m, k), device=x.device, dtype=fp8_dtype)
            x_s = torch.empty((m, 1), device=x.device, dtype=torch.float32)
            rc = _lib.load().b200_fp8_quant_per_token(
                x.data_ptr(), x.stride(0), x_q.data_ptr(), k, x_s.data_ptr(), None, m, k, 1,
                float(eps), _lib.dtype_code(x.dtype), _lib.current_stream())
            _lib.check(rc, "b200_fp8_quant_per_token(div)")
            return x_q, x_s
        raise _lib.B200Error("per_token_group_quant_fp8: group_size must be 128 (or K)")
    groups = k // 128
    x_q = torch.empty((m, k), device=x.device, dtype=fp8_dtype)
    if column_major_scales:
        rows = (m + 3) // 4 * 4 if scale_tma_aligned else m
        x_s = torch.empty((groups, rows), device=x.device, dtype=torch.float32).t()[:m]
    else:
        x_s = torch.empty((m, groups), device=x.device, dtype=torch.float32)
    rc = _lib.load().b200_fp8_quant_group128(x.data_ptr(), x.stride(0), x_q.data_ptr(), k,
                                              x_s.data_ptr(), x_s.stride(0), x_s.stride(1), m, k,
                                              float(eps), _lib.dtype_code(x.dtype),
                                              _lib.current_stream())
    _lib.check(rc, "b200_fp8_quant_group128")
    return x_q, x_s


def fp8_scaled_mm(mat_a: torch.Tensor, mat_b: torch.Tensor, scales_a: torch.Tensor,
                  scales_b: torch.Tensor, out_dtype: torch.dtype,
                  bias: Optional[torch.Tensor] = None, comm=None) -> torch.Tensor:
    """out = (A @ B) * scales_a[:,None] * scales_b[None,:] (+ bias): A [M,K] row-major e4m3,
    B [K,N] COLUMN-major e4m3 (the `.t()` view of the [N,K] weight), as sgl_kernel.fp8_scaled_mm
    (sgl-kernel/csrc/gemm/fp8_gemm_kernel.cu:1071-1146)."""
    _check_2d_cuda(mat_a, "fp8_scaled_mm")
    if mat_b.dim() != 2 or mat_b.stride(0) != 1:
        raise _lib.B200Error("fp8_scaled_mm: mat_b must be a column-major [K, N] tensor")
    m, k = mat_a.shape
    kb, n = mat_b.shape
    if kb != k:
        raise _lib.B200Error(f"fp8_scaled_mm: K mismatch ({k} vs {kb})")
    if scales_a.numel() not in (1, m) or scales_b.numel() not in (1, n):
        raise _lib.B200Error("fp8_scaled_mm: scales must be per-token [M] / per-channel [N] (or scalar)")
    out = torch.empty((m, n), device=mat_a.device, dtype=out_dtype)
    sa = scales_a.reshape(-1).to(torch.float32)
    sb = scales_b.reshape(-1).to(torch.float32)
    if bias is not None and bias.dtype != out_dtype:
        bias = bias.to(out_dtype)
    args = (mat_a.data_ptr(), mat_a.stride(0), mat_b.data_ptr(), mat_b.stride(1), sa.data_ptr(),
            1 if sa.numel() > 1 else 0, sb.data_ptr(), 1 if sb.numel() > 1 else 0, _lib.ptr(bias),
            out.data_ptr(), out.stride(0), m, n, k, _lib.dtype_code(out_dtype),
            _lib.current_stream())
    if comm is not None:  # row-parallel: GEMM + tensor-parallel all-reduce (fused for M <= 128)
        _lib.check(_lib.load().b200_fp8_gemm_rowcol_ar(comm.handle, *args),
                   "b200_fp8_gemm_rowcol_ar")
    else:
        _lib.check(_lib.load().b200_fp8_gemm_rowcol(*args), "b200_fp8_gemm_rowcol")
    return out


def fp8_blockwise_scaled_mm(mat_a: torch.Tensor, weight: torch.Tensor, scales_a: torch.Tensor,
                            scales_b: torch.Tensor, out_dtype: torch.dtype,
                            bias: Optional[torch.Tensor] = None, comm=None) -> torch.Tensor:
    """Block-scaled GEMM: A [M,K] e4m3 with [M, K/128] scales (any strides), weight [N,K] e4m3
    (K contiguous) with [ceil(N/128), ceil(K/128)] scales -> [M,N] out_dtype
    (w8a8_block_fp8_matmul, fp8_kernel.py:774-891 / sgl_kernel.fp8_blockwise_scaled_mm)."""
    _check_2d_cuda(mat_a, "fp8_blockwise_scaled_mm")
    _check_2d_cuda(weight, "fp8_blockwise_scaled_mm")
    m, k = mat_a.shape
    n, kw = weight.shape
    if kw != k:
        raise _lib.B200Error(f"fp8_blockwise_scaled_mm: K mismatch ({k} vs {kw})")
    kb = (k + 127) // 128
    if tuple(scales_a.shape) != (m, kb) or tuple(scales_b.shape) != ((n + 127) // 128, kb):
        raise _lib.B200Error("fp8_blockwise_scaled_mm: scale shapes must be [M, K/128] and "
                             "[N/128, K/128]")
    scales_a = scales_a.to(torch.float32)
    scales_b = scales_b.to(torch.float32).contiguous()
    out = torch.empty((m, n), device=mat_a.device, dtype=out_dtype)
    if bias is not None and bias.dtype != out_dtype:
        bias = bias.to(out_dtype)
    args = (mat_a.data_ptr(), mat_a.stride(0), weight.data_ptr(), weight.stride(0),
            scales_a.data_ptr(), scales_a.stride(0), scales_a.stride(1), scales_b.data_ptr(),
            scales_b.stride(0), _lib.ptr(bias), out.data_ptr(), out.stride(0), m, n, k,
            _lib.dtype_code(out_dtype), _lib.current_stream())
    if comm is not None:
        _lib.check(_lib.load().b200_fp8_gemm_block128_ar(comm.handle, *args),
                   "b200_fp8_gemm_block128_ar")
    else:
        _lib.check(_lib.load().b200_fp8_gemm_block128(*args), "b200_fp8_gemm_block128")
    return out


def fp8_gemm_smallm_config(max_m: int = -1, splits: int = -1) -> None:
    """Test/sweep hook for the decode-regime GEMM dispatch (b200_fp8_gemm_smallm_config):
    max_m = 0 forces the big-tile kernels, splits > 0 fixes the cluster split-K factor, negative
    values restore the defaults."""
    _lib.check(_lib.load().b200_fp8_gemm_smallm_config(int(max_m), int(splits)),
               "b200_fp8_gemm_smallm_config")


def comm_fuse_config(max_m: int = -1) -> None:
    """Largest token count for which row-parallel GEMMs reduce in their epilogue
    (b200_comm_fuse_config; negative = measured default)."""
    _lib.check(_lib.load().b200_comm_fuse_config(int(max_m)), "b200_comm_fuse_config")
