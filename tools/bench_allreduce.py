#!/usr/bin/env python
"""All-reduce latency over the TP group: NVLink peer-memory one-shot kernel (csrc/comm.cu) vs NCCL,
and the row-parallel FP8 GEMM with the reduce fused in its epilogue vs GEMM + all-reduce.
Run under torchrun (one rank per GPU).  Every point is a CUDA-graph of `iters` back-to-back calls,
timed with CUDA events, max over ranks; rank 0 prints one JSON line per point."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch
import torch.distributed as dist

FP8 = torch.float8_e4m3fn


def timed_graph(fn, iters, dev, reps=5):
    fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        with torch.cuda.graph(g, stream=s):
            for _ in range(iters):
                fn()
    torch.cuda.current_stream().wait_stream(s)
    g.replay()
    torch.cuda.synchronize()
    best = 1e30
    for _ in range(reps):
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        g.replay()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / iters * 1e3], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        best = min(best, float(t.item()))
    return best


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=50)
    ap.add_argument("--out", default="")
    ap.add_argument("--ar-only", action="store_true", help="skip the row-parallel GEMM points")
    args = ap.parse_args()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    from sglang_security_b200 import _lib
    from sglang_security_b200.distributed import parallel_state as ps
    from sglang_security_b200.layers.quantization.fp8_kernel import fp8_scaled_mm

    ps.initialize_model_parallel(world)
    comm = ps.get_tp_custom_all_reduce()
    assert comm is not None
    rows = []
    hidden = 4096
    for tokens in [1, 2, 4, 8, 32, 64, 128, 256, 512]:
        x = torch.randn(tokens, hidden, device=dev).to(torch.bfloat16)
        out = torch.empty_like(x)
        y = x.clone()
        us_nvls = None
        if getattr(comm, "nvls", False):   # in-switch reduction vs the peer-memory kernels
            us_nvls = timed_graph(lambda: comm.all_reduce(x, out=out), args.iters, dev)
            _lib.load().b200_comm_attach_nvls(comm.handle, None)
        us_p2p = timed_graph(lambda: comm.all_reduce(x, out=out), args.iters, dev)
        if getattr(comm, "nvls", False):
            _lib.load().b200_comm_attach_nvls(comm.handle, comm._nvls_buf)
        us_nccl = timed_graph(lambda: dist.all_reduce(y), args.iters, dev)
        row = {"op": "all_reduce", "tp": world, "tokens": tokens, "bytes": x.numel() * 2,
               "p2p_oneshot_us": round(us_p2p, 2), "nccl_us": round(us_nccl, 2),
               "nvls_us": None if us_nvls is None else round(us_nvls, 2)}
        rows.append(row)
        if rank == 0:
            print(json.dumps(row), flush=True)
    # row-parallel GEMMs of Llama-3.1-8B (o_proj K = 4096/tp, down_proj K = 14336/tp), N = 4096
    g = torch.Generator(device=dev).manual_seed(rank)
    for name, k_full in ([] if args.ar_only else [("o_proj", 4096), ("down_proj", 14336)]):
        k = k_full // world
        w = (torch.randn(hidden, k, device=dev, generator=g) * 0.5).to(FP8)
        sb = torch.rand(hidden, device=dev, generator=g) * 0.01 + 0.001
        for m in [1, 8, 16, 32, 64, 128]:
            a = torch.randn(m, k, device=dev, generator=g).to(FP8)
            sa = torch.rand(m, device=dev, generator=g) * 0.01 + 0.001
            us_f = timed_graph(lambda: fp8_scaled_mm(a, w.t(), sa, sb, torch.bfloat16, comm=comm),
                               args.iters, dev)

            def unfused_p2p():
                o = fp8_scaled_mm(a, w.t(), sa, sb, torch.bfloat16)
                comm.all_reduce(o, out=o)

            def unfused_nccl():
                o = fp8_scaled_mm(a, w.t(), sa, sb, torch.bfloat16)
                dist.all_reduce(o)

            us_g = timed_graph(lambda: fp8_scaled_mm(a, w.t(), sa, sb, torch.bfloat16),
                               args.iters, dev)
            us_u = timed_graph(unfused_p2p, args.iters, dev)
            us_n = timed_graph(unfused_nccl, args.iters, dev)
            row = {"op": "row_parallel_fp8_gemm", "layer": name, "tp": world, "M": m, "N": hidden,
                   "K_per_rank": k, "fused_us": round(us_f, 2), "gemm_only_us": round(us_g, 2),
                   "gemm_plus_p2p_ar_us": round(us_u, 2), "gemm_plus_nccl_us": round(us_n, 2)}
            rows.append(row)
            if rank == 0:
                print(json.dumps(row), flush=True)
    if rank == 0 and args.out:
        json.dump(rows, open(args.out, "w"), indent=1)
    dist.barrier()
    ps.destroy_model_parallel()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
