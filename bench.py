#!/usr/bin/env python
"""bench.py — headline benchmark of the B200 hot path (contract: task prompt / DESIGN.md §6).

Metric (BASELINE.json): decode tokens/s of Llama-3.1-8B FP8 (W8A8: per-channel e4m3 weights, dynamic
per-token e4m3 activations) at 1/2/4/8 B200, plus the decode-attention kernel's HBM GB/s vs peak.

Workload: batch 64 decode at 4k context (the geometry BASELINE configs[1] fixes for the attention
roofline, run here with the whole model of configs[2]).  One "step" = ONE WHOLE-MODEL DECODE STEP,
what the reference's `bench_one_batch.py:234-263` times: vocab-parallel embedding -> 32 x
(RMSNorm->FP8 quant, qkv GEMM, RoPE + KV-pool write, paged split-KV decode attention, o_proj GEMM
[+ TP all-reduce], RMSNorm->quant, gate_up GEMM, SiLU*mul->quant, down GEMM [+ TP all-reduce]) ->
final RMSNorm -> lm_head (+ all-gather of the vocab slices) -> greedy argmax.  Random-init weights
of the real geometry, synthetic token ids, random page table (page_size 1, worst-case gather).

N > 1 (torchrun, one rank per GPU): tensor parallel by heads / columns with the
all-reduce after every row-parallel GEMM in the timed path (NVLink peer-memory kernel or NCCL, as
`RowParallelLinear` decides) — total work is fixed => "strong" scaling.

  value    : tokens/s, the step replayed from one CUDA graph with every input resident in HBM
  e2e      : tokens/s through the same public call path with the step's host inputs (token ids,
             positions, seq_lens, out_cache_loc: what the scheduler hands the model runner each
             step) copied from pinned host memory and the sampled ids copied back and awaited
             inside the timed region; the next step's ids are the previous step's output and the
             context grows by one token per step (a real decode loop)
  roofline : dominant kernel attn_decode_mma_kernel — algorithmic bytes per launch / CUDA-event
             time per launch (library timing registry, same stream) vs the measured HBM peak
  roofline_step : (weights + KV rows + lm_head bytes of one step) / ms_per_step vs the same peak
  cpu_baseline / --impl reference : the reference's own CPU kernels compiled from its sources
             (oracle/_ref: sgl-kernel/csrc/cpu/decode.cpp attention + gemm_fp8.cpp
             fp8_scaled_mm_cpu for the four linear layers) on all host cores, ONE of the 32 layers
             at the full batch / context, scaled x32 (norms, RoPE, lm_head left out: < 2 % of it).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

WORKLOAD = dict(name="llama3.1-8b_fp8_w8a8_decode", model="llama3.1-8b", quant="fp8", bs=64,
                ctx=4096)
METRIC = "decode_tokens_per_s"
UNIT = "tokens/s"


# --------------------------------------------------------------------------------------------
def split_kernel_bytes(bs, ctx, hq, hkv, d, kv_elt=2, elt=2):
    """Algorithmic bytes of ONE attn_decode split-kernel launch: K+V rows, q, o, slot indices
    (SURVEY §8d cfg 2)."""
    return bs * ctx * hkv * 2 * d * kv_elt + bs * hq * d * elt * 2 + bs * ctx * 4


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons during the timed region."""
    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.FIELDS}",
                 "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, smax, reasons, power = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.lines:
            parts = [p.strip() for p in line.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0]))
                smax.append(float(parts[1]))
                power.append(float(parts[2]))
            except ValueError:
                continue
            for name, val in zip(names, parts[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None,
                "sm_max_mhz": max(smax) if smax else None,
                "power_w_max": max(power) if power else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        try:
            d = json.load(open(path))
            return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def ncu_traffic_per_launch(kernel_substr: str):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the dominant kernel from the newest
    `ncu --set full` capture of THIS command summarised under profiles/ (tools/ncu_summarize.py);
    None when no capture of that kernel is committed.  (A profiler cannot run inside the timed
    bench, so this field is the committed capture's figure, with its source named.)"""
    import glob
    import re

    unit = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}

    def parse(v):
        m = re.match(r"([0-9.]+)\s*(\w+)", v or "")
        return float(m.group(1)) * unit.get(m.group(2), float("nan")) if m else float("nan")

    for path in sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_ncu_decode_full_summary.json")),
                       reverse=True):
        try:
            recs = [r for r in json.load(open(path)) if kernel_substr in (r.get("Kernel Name") or "")]
            vals = [parse(r["dram__bytes_read.sum"]) + parse(r["dram__bytes_write.sum"]) for r in recs]
            if vals:
                return sum(vals) / len(vals), os.path.relpath(path, ROOT)
        except Exception:
            continue
    return None, None


# --------------------------------------------------------------------------------------------
def cpu_reference_leg(steps: int, warmup: int):
    """The reference's CPU implementation of the path, timed on the host cores on a bounded sample:
    ONE decoder layer's attention (decode_attention_cpu over the full 64 x 4096 page table) and its
    four FP8 linear layers (fp8_scaled_mm_cpu, W8A16 block-FP8 — the reference's CPU FP8 linear —
    on the Llama-3.1-8B shapes at M = 64); the other 31 layers repeat the same work on other
    buffers.  Returns (tokens/s of the whole 32-layer step, seconds per layer, info)."""
    import torch

    from oracle import ref_kernels
    from sglang_security_b200.model_executor.decode_step import GEOMETRIES

    g = GEOMETRIES[WORKLOAD["model"]]
    bs, ctx = WORKLOAD["bs"], WORKLOAD["ctx"]
    hq, hkv, d = g.heads, g.kv_heads, g.head_dim
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    gen = torch.Generator().manual_seed(0)
    size = bs * ctx
    k_cache = torch.randn((size + 1, hkv, d), generator=gen, dtype=torch.float32).to(torch.bfloat16)
    v_cache = torch.randn((size + 1, hkv, d), generator=gen, dtype=torch.float32).to(torch.bfloat16)
    req_to_token = (torch.randperm(size, generator=gen) + 1).to(torch.int32).view(bs, ctx)
    q = torch.randn((bs, hq, d), generator=gen).to(torch.bfloat16)
    k_new = torch.randn((bs, hkv, d), generator=gen).to(torch.bfloat16)
    v_new = torch.randn((bs, hkv, d), generator=gen).to(torch.bfloat16)
    req_pool_indices = torch.arange(bs, dtype=torch.int64)
    seq_lens = torch.full((bs,), ctx, dtype=torch.int64)
    loc = req_to_token[:, ctx - 1].to(torch.int64).contiguous()
    scale = d ** -0.5
    shapes = [((hq + 2 * hkv) * d, g.hidden), (g.hidden, hq * d), (2 * g.inter, g.hidden),
              (g.hidden, g.inter)]

    if ref_kernels.available():
        kind = "reference"
        desc = (f"sgl-kernel CPU decode_attention_cpu + 4 x fp8_scaled_mm_cpu "
                f"({ref_kernels.variant()} build)")
        lin = []
        for n, k in shapes:
            w = (torch.randn(n, k, generator=gen) * 0.5).to(torch.float8_e4m3fn)
            s = torch.rand((n + 127) // 128, (k + 127) // 128, generator=gen) * 0.02 + 0.005
            x = torch.randn(bs, k, generator=gen).to(torch.bfloat16)
            lin.append((x, ref_kernels.fp8_linear_prepack(w), s))

        def run():
            ref_kernels.decode_attention_ref(q, k_cache, v_cache, k_new, v_new, loc, req_to_token,
                                             req_pool_indices, seq_lens, scale)
            for x, pw, s in lin:
                ref_kernels.fp8_scaled_mm_ref(x, pw, s)
    else:
        from oracle import fp8 as OF
        from oracle.attention import decode_attention_oracle

        kind = "port"
        desc = "oracle/ restatement of torch_native decode + torch_scaled_mm FP8 linear (torch CPU)"
        lin = []
        for n, k in shapes:
            w = (torch.randn(n, k, generator=gen) * 0.5).to(torch.float8_e4m3fn)
            s = torch.rand(n, 1, generator=gen) * 0.02 + 0.005
            x = torch.randn(bs, k, generator=gen).to(torch.bfloat16)
            lin.append((x, w, s))

        def run():
            k_cache[loc] = k_new
            v_cache[loc] = v_new
            decode_attention_oracle(q, k_cache, v_cache, req_to_token, req_pool_indices, seq_lens,
                                    scale)
            for x, w, s in lin:
                OF.fp8_linear_oracle(x, w, s, None)

    for _ in range(warmup):
        run()
    times = []
    for _ in range(steps):
        t0 = time.perf_counter()
        run()
        times.append(time.perf_counter() - t0)
    t_layer = statistics.median(times)
    value = bs / (t_layer * g.layers)
    info = {"value": value, "unit": UNIT, "cores": cores, "kind": kind,
            "sample": f"1 of {g.layers} decoder layers (attention over the full bs {bs} x ctx {ctx} "
                      f"page table + qkv/o/gate_up/down FP8 linear at M={bs}; {desc}), median of "
                      f"{steps} runs = {t_layer * 1e3:.1f} ms/layer, scaled x{g.layers}"}
    return value, t_layer, info


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from sglang_security_b200.model_executor.decode_step import GEOMETRIES

    g = GEOMETRIES[WORKLOAD["model"]]
    steps = max(1, min(args.steps, 5))
    warmup = max(1, min(args.warmup, 2))
    value, t_layer, info = cpu_reference_leg(steps, warmup)
    w = WORKLOAD
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT,
            "n_gpus": args.gpus, "steps": steps, "warmup": warmup,
            "ms_per_step": t_layer * g.layers * 1e3, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "fp8_e4m3", "data": "synthetic",
            "config": workload_config(g, w["bs"], w["ctx"], g.layers, 1, "bf16",
                                      {"device": "host CPU"}),
            "cpu_baseline": info,
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0,
                    "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def workload_config(g, bs, ctx, layers, tp, kv_dtype, extra=None):
    cfg = {"workload": WORKLOAD["name"], "batch": bs, "context": ctx, "layers": layers,
           "weights": "fp8_e4m3 per-channel (W8A8, dynamic per-token activations)",
           "kv_cache": kv_dtype, "parallelism": f"tp{tp}"}
    cfg.update(extra or {})
    return cfg


# --------------------------------------------------------------------------------------------
def run_b200(args):
    import torch
    import torch.distributed as dist

    from sglang_security_b200 import _lib
    from sglang_security_b200.distributed.parallel_state import (destroy_model_parallel,
                                                                 get_tp_custom_all_reduce,
                                                                 initialize_model_parallel)
    from sglang_security_b200.model_executor.decode_step import GEOMETRIES, B200DecodeStep

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the hot path has no CPU fallback)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    initialize_model_parallel(world)
    lib = _lib.load()

    g = GEOMETRIES[WORKLOAD["model"]]
    bs, ctx = args.batch or WORKLOAD["bs"], args.context or WORKLOAD["ctx"]
    tp = world
    kv_dtype = torch.float8_e4m3fn if args.kv_dtype == "fp8" else torch.bfloat16
    total_steps = args.steps + args.warmup + 8          # context growth of the e2e decode loop
    model = B200DecodeStep(g, WORKLOAD["quant"], dev, tp_rank=rank, tp_size=tp, max_bs=bs,
                           max_ctx=ctx + total_steps, pool_tokens=bs * (ctx + total_steps),
                           kv_dtype=kv_dtype, layers=args.layers, num_kv_splits=args.splits)
    layers, hq, hkv, d = model.layers_n, model.hq, model.hkv, g.head_dim
    backend = model.attn_backend

    # page table: a random permutation of the pool (page_size 1 fragmentation), same on every rank
    gen = torch.Generator(device="cpu").manual_seed(0)
    width = ctx + total_steps
    if args.slots == "random":
        slots = (torch.randperm(bs * width, generator=gen) + 1).to(torch.int32).view(bs, width)
    else:
        slots = torch.arange(1, bs * width + 1, dtype=torch.int32).view(bs, width)
    model.req_to_token_pool.req_to_token[:bs, :width] = slots.to(dev)

    # per-step host inputs (pinned) and their device residents
    h_ids = torch.randint(0, g.vocab, (bs,), generator=gen).pin_memory()
    h_pos = torch.full((bs,), ctx - 1, dtype=torch.int64).pin_memory()
    h_lens = torch.full((bs,), ctx, dtype=torch.int64).pin_memory()
    h_loc = slots[:, ctx - 1].to(torch.int64).contiguous().pin_memory()
    h_out = torch.zeros((bs,), dtype=torch.int64).pin_memory()
    d_ids, d_pos, d_lens, d_loc = h_ids.to(dev), h_pos.to(dev), h_lens.to(dev), h_loc.to(dev)
    d_req = torch.arange(bs, dtype=torch.int64, device=dev)
    d_out = torch.zeros((bs,), dtype=torch.int64, device=dev)
    fb = model.make_forward_batch(bs, d_req, d_lens, d_loc, bs * ctx)

    def step():
        model.forward(d_ids, d_pos, fb, d_out)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def max_over_ranks(ms: float) -> float:
        if world == 1:
            return ms
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- warm-up (eager) + launch accounting -------------------------------------------------
    backend.init_forward_metadata(fb)
    n0 = lib.b200_launch_count()
    step()
    torch.cuda.synchronize()
    launches_per_step = lib.b200_launch_count() - n0
    for _ in range(2):
        step()
    torch.cuda.synchronize()
    num_splits = backend.forward_metadata.num_splits

    # ---- per-kernel split of one eager step (CUDA events around every library launch) ---------
    barrier()
    _lib.timing_enable(True)
    for _ in range(3):
        step()
    torch.cuda.synchronize()
    _lib.timing_enable(False)
    kernel_names = ("attn_decode_split", "attn_decode_fp8kv", "attn_merge", "attn_merge_quant", "dense_gemm",
                    "rope_kv_write", "kv_write",
                    "kv_write_fp8", "rmsnorm", "silu_and_mul", "fp8_quant", "fp8_gemm_smallm",
                    "fp8_gemm", "all_reduce_ll", "all_reduce_oneshot", "all_reduce_nvls")
    split = {}
    for name in kernel_names:
        n, ms = _lib.timing_read(name)
        if n:
            split[name] = {"launches_per_step": n // 3, "us_per_launch": round(ms * 1e3 / n, 2),
                           "us_per_step": round(ms * 1e3 / 3, 1)}
    attn_name = "attn_decode_fp8kv" if args.kv_dtype == "fp8" else "attn_decode_split"
    n_split = split.get(attn_name, {}).get("launches_per_step", 0) * 3
    t_split = split.get(attn_name, {}).get("us_per_launch", float("nan")) * 1e-6

    # ---- the decode step as ONE CUDA graph ----------------------------------------------------
    graph = None
    if not args.no_graph:
        md = backend.forward_metadata
        graph = torch.cuda.CUDAGraph()
        cap = torch.cuda.Stream()
        cap.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(cap):
            with torch.cuda.graph(graph, stream=cap):
                step()
        torch.cuda.current_stream().wait_stream(cap)
        backend.forward_metadata = md

    def run_step():
        if graph is not None:
            graph.replay()
        else:
            step()

    for _ in range(args.warmup):
        run_step()

    # ---- value: device-resident inputs ----------------------------------------------------------
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record()
    for _ in range(args.steps):
        run_step()
    ev1.record()
    barrier()
    ms_per_step = max_over_ranks(ev0.elapsed_time(ev1)) / args.steps
    value = bs / (ms_per_step * 1e-3)

    # ---- e2e: a real decode loop fed from host memory -----------------------------------------
    # Every step: H2D of (ids, positions, seq_lens, out_cache_loc) from pinned memory, the graph,
    # D2H of the sampled ids, and the host WAITS for them — they are the next step's input ids.
    h2d_bytes = (h_ids.numel() + h_pos.numel() + h_lens.numel() + h_loc.numel()) * 8
    d2h_bytes = h_out.numel() * 8
    col = slots.to(torch.int64)

    def e2e_step(i):
        pos = ctx - 1 + i                      # context grows by one token per step
        h_pos.fill_(pos)
        h_lens.fill_(pos + 1)
        h_loc.copy_(col[:, pos])
        d_ids.copy_(h_ids, non_blocking=True)
        d_pos.copy_(h_pos, non_blocking=True)
        d_lens.copy_(h_lens, non_blocking=True)
        d_loc.copy_(h_loc, non_blocking=True)
        run_step()
        h_out.copy_(d_out, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        h_ids.copy_(h_out)

    for i in range(min(3, args.warmup)):
        e2e_step(i)
    barrier()
    ev0.record()
    for i in range(args.steps):
        e2e_step(3 + i)
    ev1.record()
    barrier()
    e2e_ms = max_over_ranks(ev0.elapsed_time(ev1)) / args.steps
    e2e_value = bs / (e2e_ms * 1e-3)
    clocks = sampler.stop() if rank == 0 else None
    tokens_checksum = int(h_out.sum().item())      # all ranks must have sampled the same ids
    if world > 1:
        t = torch.tensor([tokens_checksum], dtype=torch.int64, device=dev)
        lo, hi = t.clone(), t.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        ranks_agree = bool(lo.item() == hi.item())
    else:
        ranks_agree = True

    comm = get_tp_custom_all_reduce()
    comm_info = None
    if world > 1:
        comm_info = {"peer_memory_all_reduce": comm is not None,
                     "crossover_bytes": getattr(comm, "crossover", None),
                     "nvls": bool(getattr(comm, "nvls", False)),
                     "row_parallel_message_bytes": bs * g.hidden * 2,
                     "all_reduces_per_step": 2 * layers + 1}
    def teardown():
        # A live CUDA graph that captured NCCL kernels keeps the communicator busy:
        # destroy_process_group() then never returns (seen at N = 2: both ranks printed / finished
        # and hung here).  Drop the graph first, then the communicators.
        nonlocal graph
        # belt and braces: the result line is out by now, so a teardown that still does not return
        # within 45 s must not hold the launcher (and every other rank) forever
        watchdog = threading.Timer(45.0, lambda: os._exit(0))
        watchdog.daemon = True
        watchdog.start()
        graph = None
        import gc

        gc.collect()
        torch.cuda.synchronize()
        if world > 1:
            destroy_model_parallel()
            dist.destroy_process_group()
        watchdog.cancel()

    if rank != 0:
        teardown()
        return

    peak, peak_src = measured_peaks()
    kv_elt = 1 if args.kv_dtype == "fp8" else 2
    per_launch_bytes = split_kernel_bytes(bs, ctx, hq, hkv, d, kv_elt)
    achieved = per_launch_bytes / t_split / 1e9 if n_split else None
    kname = "attn_decode_fp8kv_kernel" if args.kv_dtype == "fp8" else "attn_decode_mma_kernel"
    default_shape = (world == 1 and not args.layers and not args.batch and not args.context
                     and args.kv_dtype == "bf16")
    traffic, traffic_src = ncu_traffic_per_launch(kname) if default_shape else (None, None)
    roofline = {"bound": "hbm", "kernel": kname, "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": (achieved / peak) if achieved else None,
                "frac_of_nominal_8TBs": (achieved / 8000.0) if achieved else None,
                "peak_source": peak_src, "traffic": traffic, "traffic_source": traffic_src,
                "bytes_per_launch": per_launch_bytes,
                "us_per_launch": t_split * 1e6 if n_split else None, "launches_timed": n_split,
                "share_of_step": (split.get(attn_name, {}).get("us_per_step", 0.0) * 1e-3
                                  / ms_per_step) if n_split else None}
    step_bytes = model.step_algorithmic_bytes(bs, ctx)
    step_gbs = step_bytes / (ms_per_step * 1e-3) / 1e9
    roofline_step = {"bound": "hbm", "achieved": step_gbs, "peak": peak, "unit": "GB/s",
                     "frac": step_gbs / peak, "bytes_per_step_per_rank": step_bytes,
                     "weights_bytes": model.linear_weight_bytes(),
                     "lm_head_bytes": model.lm_head.numel() * 2,
                     "kv_bytes": bs * ctx * model.kv_bytes_per_token(),
                     "floor_ms": step_bytes / (peak * 1e9) * 1e3}

    cpu_baseline = None
    if world == 1 and not args.no_cpu_baseline:
        try:
            _, _, cpu_baseline = cpu_reference_leg(3, 1)
        except Exception as e:  # the GPU result must still be printed
            cpu_baseline = {"value": None, "unit": UNIT, "cores": os.cpu_count(), "kind": "port",
                            "sample": f"failed: {type(e).__name__}: {e}"}

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "fp8_e4m3", "data": "synthetic",
            "config": workload_config(g, bs, ctx, layers, tp, args.kv_dtype, {
                "heads_per_rank": f"Hq{hq}/Hkv{hkv}/D{d}", "slots": args.slots, "page_size": 1,
                "num_kv_splits": num_splits, "cuda_graph": graph is not None,
                "lm_head": "bf16 vocab-parallel + all-gather, greedy argmax",
                "l2": "inputs larger than L2: every layer streams its own "
                      f"{model.linear_weight_bytes() // layers / 1e6:.0f} MB of weights and "
                      f"{bs * ctx * model.kv_bytes_per_token() // layers / 1e6:.0f} MB of KV rows"}),
            "e2e": {"value": e2e_value, "unit": UNIT, "ms_per_step": e2e_ms,
                    "h2d_bytes_per_step": h2d_bytes, "d2h_bytes_per_step": d2h_bytes,
                    "ranks_sampled_identical_ids": ranks_agree},
            "gpu_launches": int(launches_per_step * args.steps),
            "launches_per_step": int(launches_per_step),
            "clocks": clocks, "roofline": roofline, "roofline_step": roofline_step,
            "kernel_split_eager": split, "comm": comm_info, "cpu_baseline": cpu_baseline}
    print(json.dumps(line), flush=True)
    teardown()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--splits", type=int, default=0, help="fixed KV split count (0 = heuristic)")
    ap.add_argument("--slots", default="random", choices=["random", "contiguous"])
    ap.add_argument("--kv-dtype", default="bf16", choices=["bf16", "fp8"],
                    help="KV pool dtype (bf16 = the reference's --kv-cache-dtype auto)")
    ap.add_argument("--batch", type=int, default=0, help="override the batch size (sweeps only)")
    ap.add_argument("--context", type=int, default=0, help="override the context (sweeps only)")
    ap.add_argument("--layers", type=int, default=0, help="override the layer count (debug only)")
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
