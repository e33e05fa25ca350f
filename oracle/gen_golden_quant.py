"""Generates tests/golden/{awq_dequant,fp8_quant,fp8_gemm,fp8_block}.pt by running the reference's
OWN torch-native implementations of the quantised-linear ops, whose source is extracted unmodified
(ast) from the read-only reference tree:

  awq_dequantize_torch, reverse_awq_order   sgl-kernel/tests/test_awq_dequant.py:9-57
  torch_scaled_mm                           sgl-kernel/tests/test_fp8_gemm.py:6-14
  torch_per_token_quant_fp8                 sgl-kernel/tests/test_per_token_quant_fp8.py:14-22
  native_per_token_group_quant_fp8          python/sglang/test/test_block_fp8.py:23-48
  native_static_quant_fp8                   python/sglang/test/test_block_fp8.py:97-115
  native_w8a8_block_fp8_matmul              python/sglang/test/test_block_fp8.py:276-325

These are the functions the reference's own tests compare its CUDA kernels with (the kernels
themselves cannot be built offline).  Inputs follow the reference tests' distributions
(test_fp8_gemm.py:16-25, test_block_fp8.py:369-384, test_awq_dequant.py:81-103).  Run here (needs
/root/reference); the fixtures travel to the GPU box.  TEST INFRASTRUCTURE ONLY.
"""
import ast
import os
import sys

import torch

REF = os.environ.get("B200_REFERENCE_ROOT", "/root/reference")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")
FP8 = torch.float8_e4m3fn


def load_funcs(path, names):
    """Compile the named top-level functions of `path` (and nothing else of that file)."""
    src = open(path).read()
    picked = [n for n in ast.parse(src).body if isinstance(n, ast.FunctionDef) and n.name in names]
    missing = set(names) - {n.name for n in picked}
    if missing:
        raise KeyError(f"{path}: {sorted(missing)}")
    for n in picked:
        n.decorator_list = []
    ns = {"torch": torch}
    exec(compile(ast.Module(picked, []), path, "exec"), ns)
    return [ns[n] for n in names]


def rand_fp8(shape, g):
    x = (torch.rand(shape, generator=g) - 0.5) * 2 * 448.0   # test_fp8_gemm.py:18-22
    return x.clamp(-448.0, 448.0).to(FP8)


def main():
    awq_dequantize_torch, _ = load_funcs(f"{REF}/sgl-kernel/tests/test_awq_dequant.py",
                                         ["awq_dequantize_torch", "reverse_awq_order"])
    (torch_scaled_mm,) = load_funcs(f"{REF}/sgl-kernel/tests/test_fp8_gemm.py", ["torch_scaled_mm"])
    (torch_per_token_quant_fp8,) = load_funcs(
        f"{REF}/sgl-kernel/tests/test_per_token_quant_fp8.py", ["torch_per_token_quant_fp8"])
    group_quant, static_quant, block_mm = load_funcs(
        f"{REF}/python/sglang/test/test_block_fp8.py",
        ["native_per_token_group_quant_fp8", "native_static_quant_fp8",
         "native_w8a8_block_fp8_matmul"])
    g = torch.Generator().manual_seed(0)
    i32 = torch.iinfo(torch.int32)

    # ---- AWQ dequant: random words incl. the sign bit, rand scales (test_awq_dequant.py:81-103)
    awq = []
    for rows, cols, group, dtype in [(256, 64, 128, torch.float16), (384, 40, 128, torch.bfloat16),
                                     (128, 16, 128, torch.float16), (256, 24, 64, torch.bfloat16)]:
        qweight = torch.randint(i32.min, i32.max, (rows, cols), generator=g, dtype=torch.int64).to(torch.int32)
        qzeros = torch.randint(i32.min, i32.max, (rows // group, cols), generator=g, dtype=torch.int64).to(torch.int32)
        scales = torch.rand((rows // group, cols * 8), generator=g).to(dtype)
        out = awq_dequantize_torch(qweight, scales, qzeros, group)
        awq.append(dict(qweight=qweight, qzeros=qzeros, scales=scales, group=group, out=out))
    torch.save(awq, os.path.join(GOLD, "awq_dequant.pt"))

    # ---- FP8 quantisers
    quant = []
    for m, k, dtype in [(37, 1024, torch.bfloat16), (5, 4096, torch.float16), (64, 896, torch.bfloat16)]:
        x = (torch.randn((m, k), generator=g) * 3).to(dtype)
        x[1].zero_()
        # per token: scale = absmax / 448 in fp32 (fp8_kernel.py:317-333); the reference twin takes
        # the scale and returns the bytes (an all-zero row has scale 0 -> skip it in the twin)
        scale = x.float().abs().amax(dim=-1) / 448.0
        safe = torch.where(scale > 0, scale, torch.ones_like(scale))
        q_tok = torch_per_token_quant_fp8(x, safe)
        xg = x.clone()
        xg[0, :128] = 0
        q_grp, s_grp = group_quant(xg.contiguous(), 128)
        s_static = torch.tensor([0.0123])
        q_static, _ = static_quant(x.contiguous(), s_static)
        quant.append(dict(x=x, token_scale=scale, q_token=q_tok, zero_rows=[1], x_group=xg,
                          q_group=q_grp, s_group=s_grp, static_scale=s_static, q_static=q_static))
    torch.save(quant, os.path.join(GOLD, "fp8_quant.pt"))

    # ---- FP8 per-token x per-channel GEMM (test_fp8_gemm.py:16-35)
    gemm = []
    for m, n, k, out_dtype, with_bias in [(1, 128, 512, torch.bfloat16, False),
                                          (33, 256, 512, torch.bfloat16, True),
                                          (200, 384, 1024, torch.float16, True),
                                          (130, 520, 1024, torch.bfloat16, False)]:
        a, b = rand_fp8((m, k), g), rand_fp8((n, k), g)
        sa = torch.randn((m,), generator=g) * 0.001
        sb = torch.randn((n,), generator=g) * 0.001
        bias = torch.randn((n,), generator=g).to(out_dtype) if with_bias else None
        out = torch_scaled_mm(a, b.t(), sa, sb, out_dtype, bias)
        gemm.append(dict(a=a, b=b, scale_a=sa, scale_b=sb, bias=bias, out=out))
    torch.save(gemm, os.path.join(GOLD, "fp8_gemm.pt"))

    # ---- block-scaled GEMM (test_block_fp8.py:328-396)
    block = []
    for m, n, k in [(1, 128, 128), (48, 512, 1024), (83, 640, 512), (200, 256, 1024)]:
        a, b = rand_fp8((m, k), g), rand_fp8((n, k), g)
        a_s = torch.rand((m, k // 128), generator=g) * 1e-2
        b_s = torch.rand((n // 128, k // 128), generator=g) * 1e-2
        out = block_mm(a, b, a_s, b_s, [128, 128], torch.bfloat16)
        block.append(dict(a=a, b=b, a_s=a_s, b_s=b_s, out=out))
    torch.save(block, os.path.join(GOLD, "fp8_block.pt"))
    for f in ("awq_dequant", "fp8_quant", "fp8_gemm", "fp8_block"):
        print(f, os.path.getsize(os.path.join(GOLD, f + ".pt")) // 1024, "KiB")


if __name__ == "__main__":
    sys.exit(main())
