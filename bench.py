#!/usr/bin/env python
"""bench.py — headline benchmark of the B200 hot path (contract: see task prompt / DESIGN.md §6).

Workload (BASELINE.json configs[1], SURVEY.md §8d cfg 2): Llama-3.1-8B geometry (Hq 32, Hkv 8,
D 128, 32 layers), bf16 paged decode, batch 64 at 4k context, attention + KV write only.
One "step" = one decode step of the attention path over all 32 layers: per layer the new token's
K/V rows are written to the pool (b200_kv_write) and split-KV paged attention runs over the
request's 4096 gathered KV rows (b200_attn_decode = split kernel + merge kernel), all through the
reference-facing plugin API (B200AttnBackend.forward via RadixAttention).

N > 1 (torchrun, one rank per GPU): tensor-parallel sharding by heads — rank r owns Hq/N query
heads and Hkv/N kv heads of every request and its own pool slice; the attention path
has no collective, total work is fixed => "strong" scaling.

  value   : tokens/s with all inputs resident in HBM, step replayed from a CUDA graph
  e2e     : tokens/s through the plugin API with the step's inputs (fused qkv activations of all
            layers + ForwardBatch index tensors) copied from pinned host memory and the outputs
            copied back to pinned host memory inside the timed region (the KV pool is the path's
            resident state, like weights, and is not a per-step input); the copies run on
            side streams layer by layer, overlapped with the attention of the neighbouring layers
  roofline: dominant kernel attn_decode_split, algorithmic bytes per launch / CUDA-event time
  cpu_baseline / --impl reference: the reference's own CPU decode kernel
            (sgl-kernel/csrc/cpu/decode.cpp, prebuilt in oracle/_ref) on the host cores, on a
            bounded sample (one of the 32 layers), scaled to the same metric.
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

WORKLOAD = dict(name="llama3.1-8b_bf16_paged_decode_attn+kvwrite", bs=64, ctx=4096, hq=32, hkv=8,
                d=128, layers=32)
METRIC = "decode_tokens_per_s"
UNIT = "tokens/s"


# --------------------------------------------------------------------------------------------
def algorithmic_bytes_per_layer(bs, ctx, hq, hkv, d, elt=2):
    """SURVEY §8d cfg 2: KV rows read + q + o + KV write + slot indices, per layer-step."""
    kv_read = bs * ctx * hkv * (d + d) * elt
    q_o = bs * hq * (d + d) * elt
    kv_write = bs * hkv * (d + d) * elt
    idx = bs * ctx * 4
    return kv_read + q_o + kv_write + idx


def split_kernel_bytes(bs, ctx, hq, hkv, d, elt=2):
    """Algorithmic bytes of ONE attn_decode_split launch (KV rows + q + slot indices + o)."""
    return bs * ctx * hkv * 2 * d * elt + bs * hq * d * elt * 2 + bs * ctx * 4


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
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the dominant kernel, from the
    newest `ncu --set full` capture summarised under profiles/ (tools/ncu_summarize.py); None when
    no capture of that kernel is committed."""
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
def cpu_reference_leg(steps: int, warmup: int, layers_total: int, full: bool):
    """Times the reference's CPU decode attention on the host cores on a bounded sample of the
    workload: ONE layer at the full batch/context (the other 31 layers repeat the same work on
    other pool buffers).  Returns (tokens_per_s for the whole 32-layer step, info dict)."""
    import torch

    from oracle import ref_kernels

    w = WORKLOAD
    bs, ctx, hq, hkv, d = w["bs"], w["ctx"], w["hq"], w["hkv"], w["d"]
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    g = torch.Generator().manual_seed(0)
    size = bs * ctx
    k_cache = torch.randn((size + 1, hkv, d), generator=g, dtype=torch.float32).to(torch.bfloat16)
    v_cache = torch.randn((size + 1, hkv, d), generator=g, dtype=torch.float32).to(torch.bfloat16)
    req_to_token = (torch.randperm(size, generator=g) + 1).to(torch.int32).view(bs, ctx)
    q = torch.randn((bs, hq, d), generator=g).to(torch.bfloat16)
    k_new = torch.randn((bs, hkv, d), generator=g).to(torch.bfloat16)
    v_new = torch.randn((bs, hkv, d), generator=g).to(torch.bfloat16)
    req_pool_indices = torch.arange(bs, dtype=torch.int64)
    seq_lens = torch.full((bs,), ctx, dtype=torch.int64)
    loc = req_to_token[:, ctx - 1].to(torch.int64).contiguous()
    scale = d ** -0.5

    if ref_kernels.available():
        kind = "reference"
        desc = f"sgl-kernel CPU decode_attention_cpu ({ref_kernels.variant()} build)"

        def run():
            return ref_kernels.decode_attention_ref(q, k_cache, v_cache, k_new, v_new, loc,
                                                    req_to_token, req_pool_indices, seq_lens, scale)
    else:
        from oracle.attention import decode_attention_oracle

        kind = "port"
        desc = "oracle/attention.py restatement of torch_native decode (torch CPU)"

        def run():
            k_cache[loc] = k_new
            v_cache[loc] = v_new
            return decode_attention_oracle(q, k_cache, v_cache, req_to_token, req_pool_indices,
                                           seq_lens, scale)

    for _ in range(warmup):
        run()
    times = []
    for _ in range(steps):
        t0 = time.perf_counter()
        run()
        times.append(time.perf_counter() - t0)
    t_layer = statistics.median(times)
    value = bs / (t_layer * layers_total)
    info = {"value": value, "unit": UNIT, "cores": cores, "kind": kind,
            "sample": f"1 of {layers_total} layers at full bs {bs} x ctx {ctx} ({desc}), median of "
                      f"{steps} runs = {t_layer * 1e3:.1f} ms/layer, scaled x{layers_total}"}
    return value, t_layer, info


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    steps = max(1, min(args.steps, 5))
    warmup = max(1, min(args.warmup, 2))
    value, t_layer, info = cpu_reference_leg(steps, warmup, WORKLOAD["layers"], True)
    w = WORKLOAD
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT,
            "n_gpus": args.gpus, "steps": steps, "warmup": warmup,
            "ms_per_step": t_layer * w["layers"] * 1e3, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": w["name"], "batch": w["bs"], "context": w["ctx"],
                       "layers": w["layers"], "heads": f"Hq{w['hq']}/Hkv{w['hkv']}/D{w['d']}",
                       "device": "host CPU"},
            "cpu_baseline": info,
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0,
                    "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------------------------
def run_b200(args):
    import torch
    import torch.distributed as dist

    from sglang_security_b200 import _lib
    from sglang_security_b200.layers.attention.b200_backend import B200AttnBackend
    from sglang_security_b200.layers.radix_attention import RadixAttention
    from sglang_security_b200.mem_cache.memory_pool import MHATokenToKVPool, ReqToTokenPool
    from sglang_security_b200.model_executor.forward_batch_info import ForwardBatch, ForwardMode
    from types import SimpleNamespace

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the hot path has no CPU fallback)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    lib = _lib.load()

    w = dict(WORKLOAD)
    if args.layers:
        w["layers"] = args.layers
    bs, ctx, d, layers = w["bs"], w["ctx"], w["d"], w["layers"]
    tp = world
    hq, hkv = w["hq"] // tp, max(1, w["hkv"] // tp)
    size = bs * ctx

    torch.manual_seed(1234 + rank)
    rtp = ReqToTokenPool(bs + 1, ctx + 4, dev)
    kvp = MHATokenToKVPool(size, 1, torch.bfloat16, hkv, d, layers, dev)
    for l in range(layers):
        kvp.k_buffer[l].normal_()
        kvp.v_buffer[l].normal_()
    g = torch.Generator(device="cpu").manual_seed(0)  # same slot map on every TP rank
    if args.slots == "random":
        slots = (torch.randperm(size, generator=g) + 1).to(torch.int32)
    else:
        slots = torch.arange(1, size + 1, dtype=torch.int32)
    rtp.req_to_token[:bs, :ctx] = slots.view(bs, ctx).to(dev)
    cfg = SimpleNamespace(num_attention_heads=w["hq"], context_len=ctx,
                          get_num_kv_heads=lambda t: max(1, w["hkv"] // t))
    mr = SimpleNamespace(device=dev, req_to_token_pool=rtp, token_to_kv_pool=kvp, model_config=cfg,
                         tp_size=tp, sliding_window_size=None,
                         server_args=SimpleNamespace(b200_attention_num_kv_splits=args.splits))
    backend = B200AttnBackend(mr)
    attn = [RadixAttention(hq, d, d ** -0.5, hkv, l) for l in range(layers)]

    qkv_w = (hq + 2 * hkv) * d
    # pinned host copies of every per-step input (e2e leg) and their device residents
    h_qkv = torch.randn((layers, bs, qkv_w), dtype=torch.float32).to(torch.bfloat16).pin_memory()
    h_req = torch.arange(bs, dtype=torch.int64).pin_memory()
    h_lens = torch.full((bs,), ctx, dtype=torch.int64).pin_memory()
    h_loc = slots.view(bs, ctx)[:, ctx - 1].to(torch.int64).contiguous().pin_memory()
    h_out = torch.empty((layers, bs, hq * d), dtype=torch.bfloat16).pin_memory()
    d_qkv = h_qkv.to(dev)
    d_req, d_lens, d_loc = h_req.to(dev), h_lens.to(dev), h_loc.to(dev)
    d_out = torch.empty((layers, bs, hq * d), dtype=torch.bfloat16, device=dev)

    fb = ForwardBatch(forward_mode=ForwardMode.DECODE, batch_size=bs, req_pool_indices=d_req,
                      seq_lens=d_lens, out_cache_loc=d_loc, seq_lens_sum=bs * ctx,
                      req_to_token_pool=rtp, token_to_kv_pool=kvp, attn_backend=backend)

    def step(copy_out: bool):
        backend.init_forward_metadata(fb)
        for l in range(layers):
            q, k, v = d_qkv[l].split([hq * d, hkv * d, hkv * d], dim=-1)
            o = attn[l](q, k, v, fb)
            if copy_out:
                d_out[l].copy_(o)

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
    n0 = lib.b200_launch_count()
    step(False)
    torch.cuda.synchronize()
    launches_per_step = lib.b200_launch_count() - n0
    for _ in range(max(0, args.warmup - 1)):
        step(False)
    torch.cuda.synchronize()
    num_splits = backend.forward_metadata.num_splits

    # ---- value: device-resident inputs, step replayed from a CUDA graph -----------------------
    graph = None
    if not args.no_graph:
        backend.init_cuda_graph_state(bs)
        backend.init_forward_metadata_capture_cuda_graph(bs, bs, d_req, d_lens, None,
                                                         ForwardMode.DECODE, None)
        if args.splits == 0:
            backend.forward_metadata.num_splits = num_splits
        graph = torch.cuda.CUDAGraph()
        cap_stream = torch.cuda.Stream()
        cap_stream.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(cap_stream):
            md = backend.forward_metadata
            with torch.cuda.graph(graph, stream=cap_stream):
                for l in range(layers):
                    q, k, v = d_qkv[l].split([hq * d, hkv * d, hkv * d], dim=-1)
                    attn[l](q, k, v, fb)
            backend.forward_metadata = md
        torch.cuda.current_stream().wait_stream(cap_stream)
        for _ in range(args.warmup):
            graph.replay()
        torch.cuda.synchronize()

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record()
    for _ in range(args.steps):
        if graph is not None:
            graph.replay()
        else:
            step(False)
    ev1.record()
    barrier()
    ms_total = max_over_ranks(ev0.elapsed_time(ev1))
    ms_per_step = ms_total / args.steps
    value = bs / (ms_per_step * 1e-3)

    # ---- roofline leg: per-kernel CUDA events around every launch of the dominant kernel ------
    _lib.timing_enable(True)
    barrier()
    for _ in range(args.steps):
        step(False)
    torch.cuda.synchronize()
    _lib.timing_enable(False)
    n_split, ms_split = _lib.timing_read("attn_decode_split")
    n_merge, ms_merge = _lib.timing_read("attn_merge")
    n_kvw, ms_kvw = _lib.timing_read("kv_write")
    clocks = sampler.stop() if rank == 0 else None

    # ---- e2e: plugin API, host buffers in, host buffers out ----------------------------------
    h2d_bytes = h_qkv.numel() * 2 + (h_req.numel() + h_lens.numel() + h_loc.numel()) * 8
    d2h_bytes = h_out.numel() * 2

    # The copies run on their own streams, layer by layer: layer l+1's activations come in and
    # layer l-1's output goes out while layer l computes.  Every copy of the step still happens
    # inside the timed region, and the step only counts as done once its last output is on the host.
    s_in, s_out = torch.cuda.Stream(), torch.cuda.Stream()
    ev_in = [torch.cuda.Event() for _ in range(layers)]
    ev_done = [torch.cuda.Event() for _ in range(layers)]

    def e2e_step():
        cur = torch.cuda.current_stream()
        s_in.wait_stream(cur)          # the previous step has consumed the device inputs
        with torch.cuda.stream(s_in):
            d_req.copy_(h_req, non_blocking=True)
            d_lens.copy_(h_lens, non_blocking=True)
            d_loc.copy_(h_loc, non_blocking=True)
            for l in range(layers):
                d_qkv[l].copy_(h_qkv[l], non_blocking=True)
                ev_in[l].record(s_in)
        cur.wait_event(ev_in[0])
        backend.init_forward_metadata(fb)
        for l in range(layers):
            cur.wait_event(ev_in[l])
            q, k, v = d_qkv[l].split([hq * d, hkv * d, hkv * d], dim=-1)
            d_out[l].copy_(attn[l](q, k, v, fb))
            ev_done[l].record(cur)
            s_out.wait_event(ev_done[l])
            with torch.cuda.stream(s_out):
                h_out[l].copy_(d_out[l], non_blocking=True)
        cur.wait_stream(s_out)         # the step's result is in host memory

    for _ in range(min(3, args.warmup)):
        e2e_step()
    barrier()
    ev0.record()
    for _ in range(args.steps):
        e2e_step()
    ev1.record()
    barrier()
    e2e_ms = max_over_ranks(ev0.elapsed_time(ev1)) / args.steps
    e2e_value = bs / (e2e_ms * 1e-3)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    peak, peak_src = measured_peaks()
    per_launch_bytes = split_kernel_bytes(bs, ctx, hq, hkv, d)
    t_split = (ms_split / n_split) * 1e-3 if n_split else float("nan")
    achieved = per_launch_bytes / t_split / 1e9 if n_split else None
    layer_bytes = algorithmic_bytes_per_layer(bs, ctx, hq, hkv, d)
    # the committed ncu capture is of the N=1 default workload; other shapes have no traffic figure
    traffic, traffic_src = ncu_traffic_per_launch("attn_decode_mma_kernel") \
        if (world == 1 and not args.layers) else (None, None)
    roofline = {"bound": "hbm", "kernel": "attn_decode_mma_kernel", "achieved": achieved,
                "peak": peak, "unit": "GB/s", "frac": (achieved / peak) if achieved else None,
                "frac_of_nominal_8TBs": (achieved / 8000.0) if achieved else None,
                "peak_source": peak_src, "traffic": traffic, "traffic_source": traffic_src,
                "bytes_per_launch": per_launch_bytes,
                "us_per_launch": t_split * 1e6 if n_split else None,
                "launches_timed": n_split,
                "merge_us_per_launch": (ms_merge / n_merge * 1e3) if n_merge else None,
                "kv_write_us_per_launch": (ms_kvw / n_kvw * 1e3) if n_kvw else None,
                "step_achieved_gbs": layer_bytes * layers / (ms_per_step * 1e-3) / 1e9}

    cpu_baseline = None
    if world == 1 and not args.no_cpu_baseline:
        try:
            _, _, cpu_baseline = cpu_reference_leg(3, 1, WORKLOAD["layers"], True)
        except Exception as e:  # the GPU result must still be printed
            cpu_baseline = {"value": None, "unit": UNIT, "cores": os.cpu_count(), "kind": "port",
                            "sample": f"failed: {type(e).__name__}: {e}"}

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": w["name"], "batch": bs, "context": ctx, "layers": layers,
                       "heads_per_rank": f"Hq{hq}/Hkv{hkv}/D{d}", "parallelism": f"tp{tp}",
                       "slots": args.slots, "page_size": 1, "num_kv_splits": num_splits,
                       "cuda_graph": graph is not None,
                       "l2": "inputs larger than L2: each layer reads its own "
                             f"{2 * size * hkv * d * 2 / 1e6:.0f} MB of KV rows"},
            "e2e": {"value": e2e_value, "unit": UNIT, "ms_per_step": e2e_ms,
                    "h2d_bytes_per_step": h2d_bytes, "d2h_bytes_per_step": d2h_bytes},
            "gpu_launches": int(launches_per_step * args.steps),
            "clocks": clocks, "roofline": roofline, "cpu_baseline": cpu_baseline}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--splits", type=int, default=0, help="fixed KV split count (0 = heuristic)")
    ap.add_argument("--slots", default="random", choices=["random", "contiguous"])
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
