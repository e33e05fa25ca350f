#!/usr/bin/env python
"""Whole-model decode step through the B200 hot path: BASELINE.json's headline metric
"decode tokens/s (Llama-3-8B FP8) at 1/2/4/8 B200" measured on random-init weights.

One step = embedding lookup -> L x B200DecoderLayer (fused RMSNorm->quant, FP8/INT4/bf16 linear,
RoPE->KV write, paged decode attention, SiLU*mul->quant; all-reduce after o_proj/down_proj when
TP > 1) -> final RMSNorm -> lm_head (in-tree 16-bit tcgen05 GEMM) -> greedy argmax,
replayed from one CUDA graph.  Run under torchrun for TP > 1 (one rank per GPU, NCCL).

Prints one JSON line per (batch, context) point: tokens/s, ms/step, the per-kernel time split of an
eager pass (library timing registry), and the HBM floor of the step
(weights + KV rows read once / measured HBM peak).
"""
import argparse
import json
import os
import sys
from types import SimpleNamespace

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch
import torch.distributed as dist
import torch.nn.functional as F

MODELS = {
    # hidden, heads, kv_heads, head_dim, inter, layers, vocab, rope_theta, qkv_bias
    "llama3.1-8b": (4096, 32, 8, 128, 14336, 32, 128256, 500000.0, False),
    "llama3.1-70b": (8192, 64, 8, 128, 28672, 80, 128256, 500000.0, False),
    "qwen2.5-32b": (5120, 40, 8, 128, 27648, 64, 152064, 1000000.0, True),
    "qwen2.5-0.5b": (896, 14, 2, 64, 4864, 24, 151936, 1000000.0, True),
}
FP8 = torch.float8_e4m3fn


def quant_config(kind):
    from sglang_security_b200.layers.quantization.awq import AWQConfig
    from sglang_security_b200.layers.quantization.fp8 import Fp8Config
    from sglang_security_b200.layers.quantization.w8a8_fp8 import W8A8Fp8Config

    if kind == "bf16":
        return None
    if kind == "fp8":            # per-channel weights, dynamic per-token activations
        return W8A8Fp8Config(True)
    if kind == "fp8_block":      # 128x128 weight blocks, 1x128 activation groups
        return Fp8Config(True, "dynamic", weight_block_size=[128, 128])
    if kind == "awq":
        return AWQConfig(4, 128, True)
    raise ValueError(kind)


def fill_random(layer_mod, kind, gen_seed):
    """Random-init weights of the right storage type, then process_weights_after_loading."""
    g = torch.Generator(device="cuda").manual_seed(gen_seed)
    for name in ("qkv_proj", "o_proj", "gate_up_proj", "down_proj"):
        lin = getattr(layer_mod, name)
        for pname, p in lin.named_parameters(recurse=False):
            if p.dtype == FP8:
                w = torch.randn(p.shape, device="cuda", generator=g) * 0.5
                p.data.copy_(w.to(FP8))
            elif p.dtype == torch.int32:
                p.data.copy_(torch.randint(-2 ** 31, 2 ** 31 - 1, p.shape, device="cuda",
                                           generator=g, dtype=torch.int64).to(torch.int32))
            elif "scale" in pname:
                p.data.copy_(torch.rand(p.shape, device="cuda", generator=g) * 0.02 + 0.005)
            elif pname == "bias":
                p.data.normal_(0, 0.02, generator=g)
            else:
                p.data.normal_(0, 0.02, generator=g)
        lin.quant_method.process_weights_after_loading(lin)
    for n in ("input_layernorm", "post_attention_layernorm"):
        getattr(layer_mod, n).weight.data = torch.ones_like(getattr(layer_mod, n).weight.data).to(
            torch.bfloat16)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="llama3.1-8b", choices=list(MODELS))
    ap.add_argument("--quant", default="fp8", choices=["fp8", "fp8_block", "bf16", "awq"])
    ap.add_argument("--bs", default="1,8,32,64,128,256")
    ap.add_argument("--ctx", default="1024,4096")
    ap.add_argument("--layers", type=int, default=0)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--no-fuse", action="store_true")
    ap.add_argument("--pdl", type=int, default=1, help="programmatic dependent launch on/off")
    ap.add_argument("--max-kv-gb", type=float, default=80.0)
    ap.add_argument("--kv-dtype", default="bf16", choices=["bf16", "fp8"],
                    help="KV pool dtype (fp8 = float8_e4m3fn, --kv-cache-dtype fp8_e4m3)")
    ap.add_argument("--out", default="")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    from sglang_security_b200 import _lib
    from sglang_security_b200.distributed.parallel_state import initialize_model_parallel
    from sglang_security_b200.layers.attention.b200_backend import B200AttnBackend
    from sglang_security_b200.layers.dense_kernel import argmax_rows, gemm_16bit
    from sglang_security_b200.layers.layernorm import RMSNorm
    from sglang_security_b200.layers.rotary_embedding import RotaryEmbedding
    from sglang_security_b200.mem_cache.memory_pool import MHATokenToKVPool, ReqToTokenPool
    from sglang_security_b200.model_executor.decoder_layer import B200DecoderLayer
    from sglang_security_b200.model_executor.forward_batch_info import ForwardBatch, ForwardMode

    initialize_model_parallel(world)
    _lib.set_pdl(args.pdl)
    hidden, heads, kv_heads, d, inter, layers, vocab, theta, qkv_bias = MODELS[args.model]
    if args.layers:
        layers = args.layers
    tp = world
    hq, hkv = heads // tp, max(1, kv_heads // tp)
    bss = [int(x) for x in args.bs.split(",")]
    ctxs = [int(x) for x in args.ctx.split(",")]
    # KV pool sized for the largest (bs, ctx) point that fits the budget
    kv_fp8 = args.kv_dtype == "fp8"
    per_tok = 2 * hkv * d * (1 if kv_fp8 else 2) * layers
    points = [(b, c) for c in ctxs for b in bss if b * c * per_tok <= args.max_kv_gb * 1e9]
    pool_tokens = max(b * c for b, c in points)
    max_bs, max_ctx = max(b for b, _ in points), max(c for _, c in points)

    torch.manual_seed(1234)
    qc = quant_config(args.quant)
    rope = RotaryEmbedding(d, d, max_ctx + 8, theta, True, torch.bfloat16).to(dev)
    rtp = ReqToTokenPool(max_bs + 1, max_ctx + 8, dev)
    kvp = MHATokenToKVPool(pool_tokens + max_bs, 1, FP8 if kv_fp8 else torch.bfloat16, hkv, d,
                           layers, dev)
    for l in range(layers):
        for buf in (kvp.k_buffer[l], kvp.v_buffer[l]):
            if kv_fp8:  # random e4m3 bytes without the NaN patterns
                buf.random_(0, 256)
                buf[(buf & 0x7f) == 0x7f] = 0x38
            else:
                buf.normal_()
    cfg = SimpleNamespace(num_attention_heads=heads, context_len=max_ctx + 4,
                          get_num_kv_heads=lambda t: max(1, kv_heads // t))
    backend = B200AttnBackend(SimpleNamespace(
        device=dev, req_to_token_pool=rtp, token_to_kv_pool=kvp, model_config=cfg, tp_size=tp,
        server_args=SimpleNamespace()))
    mods = []
    for l in range(layers):
        L = B200DecoderLayer(hidden, heads, kv_heads, d, inter, l, rope, quant_config=qc,
                             params_dtype=torch.bfloat16, rms_norm_eps=1e-5, qkv_bias=qkv_bias,
                             fuse=not args.no_fuse, tp_rank=rank, tp_size=tp).to(dev)
        fill_random(L, args.quant, 100 + l)
        mods.append(L)
    embed = (torch.randn(vocab, hidden, device=dev) * 0.02).to(torch.bfloat16)
    lm_head = (torch.randn(vocab // tp, hidden, device=dev) * 0.02).to(torch.bfloat16)
    final_norm = RMSNorm(hidden, 1e-5).to(dev)
    final_norm.weight.data = final_norm.weight.data.to(torch.bfloat16)

    lin_bytes = 0
    for L in mods:
        for n in ("qkv_proj", "o_proj", "gate_up_proj", "down_proj"):
            for p in getattr(L, n).parameters():
                lin_bytes += p.numel() * p.element_size()
    peak = 6548.5
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        pass

    g = torch.Generator().manual_seed(0)
    rows = []
    for bs, ctx in points:
        slots = (torch.randperm(bs * ctx, generator=g) + 1).to(torch.int32).view(bs, ctx)
        rtp.req_to_token[:bs, :ctx] = slots.to(dev)
        ids = torch.randint(0, vocab, (bs,), generator=g).to(dev)
        positions = torch.full((bs,), ctx - 1, dtype=torch.int64, device=dev)
        fb = ForwardBatch(forward_mode=ForwardMode.DECODE, batch_size=bs,
                          req_pool_indices=torch.arange(bs, dtype=torch.int64, device=dev),
                          seq_lens=torch.full((bs,), ctx, dtype=torch.int64, device=dev),
                          out_cache_loc=slots[:, ctx - 1].to(torch.int64).to(dev),
                          seq_lens_sum=bs * ctx, req_to_token_pool=rtp, token_to_kv_pool=kvp,
                          attn_backend=backend)
        out_tok = torch.empty((bs,), dtype=torch.int64, device=dev)

        def step():
            h = embed[ids]
            residual = None
            for L in mods:
                h, residual = L(positions, h, fb, residual)
            h, _ = final_norm(h, residual)
            logits = gemm_16bit(h, lm_head)
            argmax_rows(logits, out_tok)

        backend.init_forward_metadata(fb)
        n0 = _lib.load().b200_launch_count()
        step()
        torch.cuda.synchronize()
        launches = _lib.load().b200_launch_count() - n0
        step()
        # per-kernel split (eager, CUDA events around every library launch)
        _lib.timing_enable(True)
        step()
        torch.cuda.synchronize()
        _lib.timing_enable(False)
        split = {}
        for name in ("rmsnorm", "fp8_gemm_smallm", "fp8_block_gemm_smallm", "fp8_gemm",
                     "fp8_block_gemm", "w4a16_gemm", "w4a16_gemv", "rope_kv_write",
                     "attn_decode_split", "attn_decode_fp8kv", "kv_write_fp8", "attn_merge",
                     "attn_merge_quant", "w4a16_gemv2", "dense_gemm", "all_reduce_ll",
                     "all_reduce_oneshot", "fp8_quant", "silu_and_mul", "kv_write"):
            n, ms = _lib.timing_read(name)
            if n:
                split[name] = {"launches": n, "us_total": round(ms * 1e3, 1)}
        num_splits = backend.forward_metadata.num_splits
        md = backend.forward_metadata
        graph = torch.cuda.CUDAGraph()
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            with torch.cuda.graph(graph, stream=s):
                step()
        torch.cuda.current_stream().wait_stream(s)
        backend.forward_metadata = md
        for _ in range(3):
            graph.replay()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.steps):
            graph.replay()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / args.steps
        if world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        kv_bytes = bs * ctx * per_tok
        floor_ms = (lin_bytes + kv_bytes + lm_head.numel() * 2) / (peak * 1e9) * 1e3
        row = {"model": args.model, "quant": args.quant, "kv_dtype": args.kv_dtype, "tp": tp, "bs": bs, "ctx": ctx,
               "layers": layers, "fused": not args.no_fuse, "pdl": args.pdl, "ms_per_step": round(ms, 4),
               "tokens_per_s": round(bs / ms * 1e3, 1), "hbm_floor_ms": round(floor_ms, 4),
               "frac_of_hbm_floor": round(floor_ms / ms, 4), "num_kv_splits": num_splits,
               "b200_launches_per_step": int(launches), "kernel_split_eager": split}
        if rank == 0:
            print(json.dumps(row), flush=True)
        rows.append(row)
        del graph
    if rank == 0 and args.out:
        os.makedirs(os.path.dirname(os.path.abspath(args.out)), exist_ok=True)
        json.dump(rows, open(args.out, "w"), indent=1)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
