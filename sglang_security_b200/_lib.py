"""ctypes binding of libb200hotpath.so (the C-ABI declared in include/b200_hotpath.h).

There is deliberately NO fallback: if the shared library is missing or a call returns a
non-zero status this module raises.  The product path never routes through a CPU or
PyTorch-eager substitute.
"""
from __future__ import annotations

import ctypes
from ctypes import c_char_p, c_float, c_int, c_int64, c_size_t, c_void_p
from pathlib import Path

_LIB_PATH = Path(__file__).resolve().parent / "lib" / "libb200hotpath.so"
_lib = None

# dtype codes (enum b200_dtype)
BF16, F16, F32, FP8_E4M3 = 0, 1, 2, 3
KV_FP8_E5M2 = 0x100  # B200_KV_FP8_E5M2: flag on the dtype argument of the FP8-KV-pool entry points

_SIGNATURES = {
    "b200_version": (c_int, []),
    "b200_last_error": (c_char_p, []),
    "b200_last_variant": (c_char_p, []),
    "b200_launch_count": (c_int64, []),
    "b200_set_pdl": (c_int, [c_int]),
    "b200_timing_enable": (c_int, [c_int]),
    "b200_timing_read": (c_int, [c_char_p, ctypes.POINTER(c_int64), ctypes.POINTER(ctypes.c_double)]),
    "b200_kv_write": (c_int, [c_void_p, c_void_p, c_int64, c_int64, c_void_p, c_void_p, c_void_p,
                              c_int, c_int, c_int, c_void_p]),
    "b200_kv_write_fp8": (c_int, [c_void_p, c_void_p, c_int64, c_int64, c_void_p, c_void_p, c_void_p,
                                  c_int, c_int, c_float, c_float, c_int, c_int, c_void_p]),
    "b200_attn_decode_fp8kv_split_align": (c_int, []),
    "b200_attn_decode_fp8kv": (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_int64, c_void_p,
                                       c_int64, c_void_p, c_int64, c_void_p, c_void_p, c_int, c_int,
                                       c_int, c_int, c_int, c_float, c_float, c_float, c_float,
                                       c_int, c_void_p, c_size_t, c_int, c_void_p]),
    "b200_kv_dequant_gather": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_int64, c_void_p, c_void_p,
                                       c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int,
                                       c_int64, c_int, c_void_p]),
    "b200_attn_extend_scaled": (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_int64, c_void_p,
                                        c_int64, c_void_p, c_int64, c_void_p, c_void_p, c_void_p,
                                        c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int,
                                        c_float, c_float, c_float, c_int, c_int, c_void_p]),
    "b200_write_req_to_token": (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_void_p,
                                        c_void_p, c_int, c_void_p]),
    "b200_get_last_loc": (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_int, c_void_p]),
    "b200_alloc_extend": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int,
                                  c_int, c_void_p]),
    "b200_alloc_decode": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int,
                                  c_void_p]),
    "b200_compute_position": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p]),
    "b200_kv_indices": (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_void_p,
                                c_void_p, c_int, c_void_p]),
    "b200_attn_decode_workspace_bytes": (c_size_t, [c_int, c_int, c_int, c_int]),
    "b200_attn_decode_num_splits": (c_int, [c_int, c_int, c_int, c_int, c_int64, c_int]),
    "b200_attn_decode_split_align": (c_int, [c_int, c_int]),
    "b200_attn_decode": (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_int64, c_void_p,
                                 c_int64, c_void_p, c_int64, c_void_p, c_void_p, c_int, c_int,
                                 c_int, c_int, c_int, c_float, c_float, c_int, c_void_p, c_size_t,
                                 c_int, c_void_p]),
    # split-KV scratch built into the library (8 MiB, B200_ATTN_DECODE_NOALLOC_SCRATCH_BYTES): allocates nothing
    "b200_attn_decode_noalloc": (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_int64, c_void_p,
                                         c_int64, c_void_p, c_int64, c_void_p, c_void_p, c_int, c_int,
                                         c_int, c_int, c_int, c_float, c_float, c_int, c_int,
                                         c_void_p]),
    "b200_attn_decode_fused": (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_int64, c_void_p,
                                       c_int64, c_void_p, c_int64, c_void_p, c_void_p, c_int, c_int,
                                       c_int, c_int, c_int, c_float, c_float, c_int, c_void_p,
                                       c_size_t, c_int, c_void_p, c_size_t, c_void_p, c_int64,
                                       c_void_p, c_void_p]),
    "b200_attn_decode_fp8kv_fused": (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_int64, c_void_p,
                                             c_int64, c_void_p, c_int64, c_void_p, c_void_p, c_int,
                                             c_int, c_int, c_int, c_int, c_float, c_float, c_float,
                                             c_float, c_int, c_void_p, c_size_t, c_int, c_void_p,
                                             c_size_t, c_void_p, c_int64, c_void_p, c_void_p]),
    "b200_attn_decode_fused_ticket_bytes": (c_size_t, [c_int, c_int, c_int]),
    "b200_attn_merge": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_void_p, c_int, c_int,
                                c_int, c_int, c_int, c_int, c_void_p]),
    "b200_attn_extend": (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_int64, c_void_p, c_int64,
                                 c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_void_p,
                                 c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_float,
                                 c_float, c_int, c_int, c_void_p]),
    "b200_fp8_quant_per_token": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_void_p,
                                         c_int, c_int, c_int, c_float, c_int, c_void_p]),
    "b200_fp8_quant_group128": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_int64,
                                        c_int64, c_int, c_int, c_float, c_int, c_void_p]),
    "b200_fp8_gemm_rowcol": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_int,
                                     c_void_p, c_int, c_void_p, c_void_p, c_int64, c_int, c_int,
                                     c_int, c_int, c_void_p]),
    "b200_fp8_gemm_smallm_config": (c_int, [c_int, c_int]),
    "b200_fp8_gemm_block128": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_int64,
                                       c_int64, c_void_p, c_int64, c_void_p, c_void_p, c_int64,
                                       c_int, c_int, c_int, c_int, c_void_p]),
    "b200_comm_handle_bytes": (c_int, []),
    "b200_comm_buffer_bytes": (c_size_t, [c_int, c_size_t]),
    "b200_comm_alloc": (c_int, [c_size_t, ctypes.POINTER(c_void_p), c_void_p]),
    "b200_comm_open": (c_int, [c_void_p, ctypes.POINTER(c_void_p)]),
    "b200_comm_close": (c_int, [c_void_p]),
    "b200_comm_free": (c_int, [c_void_p]),
    "b200_comm_create": (c_int, [c_int, c_int, ctypes.POINTER(c_void_p), c_size_t,
                                 ctypes.POINTER(c_void_p)]),
    "b200_comm_destroy": (c_int, [c_void_p]),
    "b200_comm_max_bytes": (c_size_t, [c_void_p]),
    "b200_comm_all_reduce": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int, c_void_p]),
    "b200_comm_fuse_config": (c_int, [c_int]),
    "b200_nvls_supported": (c_int, []),
    "b200_nvls_buffer_bytes": (c_size_t, [c_int, c_size_t]),
    "b200_nvls_create": (c_int, [c_int, c_size_t, ctypes.POINTER(c_void_p), ctypes.POINTER(c_int)]),
    "b200_nvls_import": (c_int, [c_int, c_int, c_size_t, ctypes.POINTER(c_void_p)]),
    "b200_nvls_add_device": (c_int, [c_void_p]),
    "b200_nvls_bind": (c_int, [c_void_p, ctypes.POINTER(c_void_p), ctypes.POINTER(c_void_p)]),
    "b200_nvls_free": (c_int, [c_void_p]),
    "b200_comm_attach_nvls": (c_int, [c_void_p, c_void_p]),
    "b200_fp8_gemm_rowcol_ar": (c_int, [c_void_p, c_void_p, c_int64, c_void_p, c_int64, c_void_p,
                                        c_int, c_void_p, c_int, c_void_p, c_void_p, c_int64, c_int,
                                        c_int, c_int, c_int, c_void_p]),
    "b200_fp8_gemm_block128_ar": (c_int, [c_void_p, c_void_p, c_int64, c_void_p, c_int64, c_void_p,
                                          c_int64, c_int64, c_void_p, c_int64, c_void_p, c_void_p,
                                          c_int64, c_int, c_int, c_int, c_int, c_void_p]),
    "b200_argmax_rows": (c_int, [c_void_p, c_int64, c_void_p, c_int, c_int, c_int, c_void_p]),
    "b200_gemm_16bit": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_void_p, c_int64,
                                c_int, c_int, c_int, c_int, c_void_p]),
    "b200_w4a16_gemv": (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_void_p,
                                c_void_p, c_int64, c_int, c_int, c_int, c_int, c_int, c_void_p]),
    "b200_w4a16_gemv_config": (c_int, [c_int, c_int]),
    "b200_w4a16_gemv_v2_repack": (c_int, [c_void_p, c_void_p, c_int, c_int, c_void_p]),
    "b200_w4a16_gemv_v2": (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_void_p,
                                   c_void_p, c_int64, c_int, c_int, c_int, c_int, c_int, c_void_p]),
    "b200_w4a16_gemv_v2_supported": (c_int, [c_int, c_int, c_int, c_int]),
    "b200_w4a16_gemv_v2_config": (c_int, [c_int]),
    "b200_rmsnorm": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_void_p, c_int64,
                             c_void_p, c_int64, c_void_p, c_int64, c_int64, c_int, c_int, c_float,
                             c_int, c_float, c_int, c_void_p]),
    "b200_silu_and_mul": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_int64,
                                  c_void_p, c_int64, c_int64, c_int, c_int, c_int, c_float, c_int,
                                  c_void_p]),
    "b200_rope_kv_write": (c_int, [c_void_p, c_void_p, c_int64, c_void_p, c_int64, c_void_p,
                                   c_int64, c_void_p, c_int, c_int, c_int, c_int, c_void_p,
                                   c_void_p, c_int64, c_void_p, c_int, c_int, c_int, c_int, c_int,
                                   c_void_p]),
    "b200_rope_kv_write_fp8": (c_int, [c_void_p, c_void_p, c_int64, c_void_p, c_int64, c_void_p,
                                       c_int64, c_void_p, c_int, c_int, c_int, c_int, c_void_p,
                                       c_void_p, c_int64, c_void_p, c_float, c_float, c_int, c_int,
                                       c_int, c_int, c_int, c_int, c_void_p]),
    "b200_awq_dequant": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int,
                                 c_int, c_void_p]),
    "b200_w4a16_unpack_zeros": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int,
                                        c_void_p]),
    "b200_w4a16_repack_gptq": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p]),
    "b200_w4a16_gemm": (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_void_p,
                                c_void_p, c_int64, c_int, c_int, c_int, c_int, c_int, c_void_p]),
}


class B200Error(RuntimeError):
    """Raised when a b200_* entry point reports a non-zero status."""


def lib_path() -> Path:
    return _LIB_PATH


def exported_symbols():
    return list(_SIGNATURES)


def load():
    """Load the shared library once; raise if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not _LIB_PATH.exists():
        raise B200Error(
            f"{_LIB_PATH} is missing: build it with `python -m sglang_security_b200.build` "
            "(there is no CPU / eager fallback for the hot path)")
    lib = ctypes.CDLL(str(_LIB_PATH))
    for name, (restype, argtypes) in _SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the .so is stale: fail loudly
        fn.restype = restype
        fn.argtypes = argtypes
    _lib = lib
    return lib


def check(status: int, what: str) -> None:
    if status != 0:
        msg = load().b200_last_error()
        raise B200Error(f"{what} failed (status {status}): {msg.decode() if msg else ''}")


def last_variant() -> str:
    """Which kernel instance the last GEMM-family call of this thread dispatched to (test hook)."""
    v = load().b200_last_variant()
    return v.decode() if v else ""


def set_pdl(on) -> int:
    """Programmatic dependent launch: False/0 off, True/1 on for every kernel, other ints = bit mask
    over kernel families (experiments).  Returns the previous mask."""
    return int(load().b200_set_pdl(int(on)))


def timing_enable(on: bool) -> None:
    load().b200_timing_enable(1 if on else 0)


def timing_read(name: str = ""):
    """-> (launches, total_ms) of the timed launches of kernel `name` since the last read."""
    n, ms = c_int64(0), ctypes.c_double(0.0)
    load().b200_timing_read(name.encode(), ctypes.byref(n), ctypes.byref(ms))
    return n.value, ms.value


def ptr(t):
    """Device pointer of a torch tensor (None -> NULL)."""
    return None if t is None else t.data_ptr()


def dtype_code(torch_dtype) -> int:
    import torch

    table = {torch.bfloat16: BF16, torch.float16: F16, torch.float32: F32,
             torch.float8_e4m3fn: FP8_E4M3}
    if torch_dtype not in table:
        raise B200Error(f"dtype {torch_dtype} is not supported by the B200 hot path")
    return table[torch_dtype]


def is_fp8_kv(torch_dtype) -> bool:
    import torch

    return torch_dtype in (torch.float8_e4m3fn, torch.float8_e5m2)


def kv_dtype_code(act_dtype, pool_dtype) -> int:
    """dtype argument of an FP8-KV-pool entry point: activation dtype code | the e5m2 flag."""
    import torch

    return dtype_code(act_dtype) | (KV_FP8_E5M2 if pool_dtype == torch.float8_e5m2 else 0)


def current_stream() -> int:
    import torch

    return torch.cuda.current_stream().cuda_stream
