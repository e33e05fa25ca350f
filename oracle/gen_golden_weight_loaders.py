"""Generates tests/golden/weight_loaders.pt by running the reference's OWN fused-shard weight loaders:
  python/sglang/srt/layers/parameter.py   imported UNMODIFIED by file path (it only needs torch):
      ModelWeightParameter, PackedvLLMParameter, GroupQuantScaleParameter, ChannelQuantScaleParameter,
      BlockQuantScaleParameter, PerTensorScaleParameter and their load_*_weight methods (:25-375)
  python/sglang/srt/layers/linear.py      methods extracted unmodified (ast) and bound to bare
      objects carrying the fields the constructors set (:299-341, :463-520, :750-836, :1130-1200):
      ColumnParallelLinear.weight_loader_v2 (:422-440), MergedColumnParallelLinear.weight_loader_v2 +
      _load_fused_module_from_checkpoint (:660-747), QKVParallelLinear.weight_loader_v2 +
      _get_shard_offset_mapping + _get_shard_size_mapping + _load_fused_module_from_checkpoint
      (:837-934), RowParallelLinear.weight_loader_v2 (:1256-1276)
One full checkpoint per case is loaded on every simulated TP rank; the fixture stores the checkpoint
tensors and the rank's resulting parameter.  Parameter shapes follow the reference's quant methods:
unquantised weight (linear.py:143-160), AWQ qweight / qzeros / scales (quantization/awq.py:139-175),
FP8 per-tensor and 128x128 block scales (quantization/fp8.py:209-300), GPTQ qweight packed along K.
Run here (needs /root/reference).  TEST INFRASTRUCTURE ONLY."""
import ast
import importlib.util
import os
import sys
import zlib
from types import SimpleNamespace
from typing import List, Optional, Tuple

import torch

REF = "/root/reference/python/sglang/srt/layers"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def load_reference():
    spec = importlib.util.spec_from_file_location("ref_parameter", f"{REF}/parameter.py")
    P = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(P)
    ns = {"torch": torch, "Optional": Optional, "List": List, "Tuple": Tuple,
          "Parameter": torch.nn.Parameter}
    for name in ("BasevLLMParameter", "PackedColumnParameter", "PackedvLLMParameter",
                 "PerTensorScaleParameter", "RowvLLMParameter", "BlockQuantScaleParameter",
                 "_ColumnvLLMParameter"):
        ns[name] = getattr(P, name)
    wanted = {"ColumnParallelLinear": ["weight_loader_v2"],
              "MergedColumnParallelLinear": ["weight_loader_v2", "_load_fused_module_from_checkpoint"],
              "QKVParallelLinear": ["weight_loader_v2", "_get_shard_offset_mapping",
                                    "_get_shard_size_mapping", "_load_fused_module_from_checkpoint"],
              "RowParallelLinear": ["weight_loader_v2"]}
    classes = {}
    path = f"{REF}/linear.py"
    for node in ast.parse(open(path).read()).body:
        if isinstance(node, ast.ClassDef) and node.name in wanted:
            fns = [n for n in node.body if isinstance(n, ast.FunctionDef) and n.name in wanted[node.name]]
            assert sorted(f.name for f in fns) == sorted(wanted[node.name]), node.name
            for f in fns:
                f.decorator_list = []
            scope = dict(ns)
            exec(compile(ast.Module(fns, []), path, "exec"), scope)
            classes[node.name] = type("Ref" + node.name, (), {f.name: scope[f.name] for f in fns})
    return P, classes


def ref_layer(cls, tp_rank, tp_size, block=None, **fields):
    obj = cls()
    obj.tp_rank, obj.tp_size, obj.use_presharded_weights = tp_rank, tp_size, False
    obj.quant_method = SimpleNamespace(quant_config=SimpleNamespace(weight_block_size=block))
    for k, v in fields.items():
        setattr(obj, k, v)
    return obj


def qkv_fields(heads, kv_heads, d, tp):
    """QKVParallelLinear.__init__ (linear.py:803-821)."""
    num_heads = heads // tp
    if tp >= kv_heads:
        num_kv, replicas = 1, tp // kv_heads
    else:
        num_kv, replicas = kv_heads // tp, 1
    return dict(total_num_heads=heads, total_num_kv_heads=kv_heads, head_size=d, num_heads=num_heads,
                num_kv_heads=num_kv, num_kv_head_replicas=replicas,
                output_sizes=[num_heads * d * tp, num_kv * d * tp, num_kv * d * tp])


# ---- reference-side parameters, shaped as the reference's quant methods shape them -------------------
def make_params(P, quant, in_pp, out_pp, n_logical, dtype):
    """-> {name: reference parameter} of one linear layer on one rank (weight_loader is unused here:
    the layer-level weight_loader_v2 is called directly, as model load_weights does)."""
    wl = None
    if quant is None:
        return {"weight": P.ModelWeightParameter(data=torch.zeros(out_pp, in_pp, dtype=dtype),
                                                 input_dim=1, output_dim=0, weight_loader=wl)}
    if quant[0] == "awq":            # quantization/awq.py:150-189
        group = quant[1]
        pk = dict(input_dim=0, output_dim=1, packed_dim=1, packed_factor=8, weight_loader=wl)
        return {"qweight": P.PackedvLLMParameter(data=torch.zeros(in_pp, out_pp // 8, dtype=torch.int32), **pk),
                "qzeros": P.PackedvLLMParameter(data=torch.zeros(in_pp // group, out_pp // 8, dtype=torch.int32), **pk),
                "scales": P.GroupQuantScaleParameter(data=torch.zeros(in_pp // group, out_pp, dtype=dtype),
                                                     input_dim=0, output_dim=1, weight_loader=wl)}
    if quant[0] == "fp8_tensor":     # quantization/fp8.py:265-309 (serialized checkpoint, static activations)
        return {"weight": P.ModelWeightParameter(data=torch.zeros(out_pp, in_pp, dtype=torch.float8_e4m3fn),
                                                 input_dim=1, output_dim=0, weight_loader=wl),
                "weight_scale": P.PerTensorScaleParameter(data=torch.zeros(n_logical, dtype=torch.float32),
                                                          weight_loader=wl),
                "input_scale": P.PerTensorScaleParameter(data=torch.zeros(n_logical, dtype=torch.float32),
                                                         weight_loader=wl)}
    if quant[0] == "fp8_block":      # quantization/fp8.py:265-292
        bn, bk = quant[1]
        return {"weight": P.ModelWeightParameter(data=torch.zeros(out_pp, in_pp, dtype=torch.float8_e4m3fn),
                                                 input_dim=1, output_dim=0, weight_loader=wl),
                "weight_scale_inv": P.BlockQuantScaleParameter(
                    data=torch.zeros((out_pp + bn - 1) // bn, (in_pp + bk - 1) // bk, dtype=torch.float32),
                    input_dim=1, output_dim=0, weight_loader=wl)}
    if quant[0] == "gptq":           # vllm gptq.py create_weights (the method the reference delegates to)
        group = quant[1]
        return {"qweight": P.PackedvLLMParameter(data=torch.zeros(in_pp // 8, out_pp, dtype=torch.int32),
                                                 input_dim=0, output_dim=1, packed_dim=0, packed_factor=8,
                                                 weight_loader=wl),
                "qzeros": P.PackedvLLMParameter(data=torch.zeros(in_pp // group, out_pp // 8, dtype=torch.int32),
                                                input_dim=0, output_dim=1, packed_dim=1, packed_factor=8,
                                                weight_loader=wl),
                "scales": P.GroupQuantScaleParameter(data=torch.zeros(in_pp // group, out_pp, dtype=dtype),
                                                     input_dim=0, output_dim=1, weight_loader=wl),
                "g_idx": P.RowvLLMParameter(data=torch.zeros(in_pp, dtype=torch.int32), input_dim=0,
                                            weight_loader=wl)}
    raise KeyError(quant)


def checkpoint_tensor(g, pname, quant, rows_out, cols_in, dtype):
    """One logical projection's tensor as it sits in a checkpoint (full, unsharded)."""
    if pname == "weight":
        w = torch.randn(rows_out, cols_in, generator=g)
        return w.to(torch.float8_e4m3fn) if quant and quant[0].startswith("fp8") else w.to(dtype)
    if pname in ("weight_scale", "input_scale"):
        return torch.rand((), generator=g) + 0.1
    if pname == "weight_scale_inv":
        bn, bk = quant[1]
        return torch.rand((rows_out + bn - 1) // bn, (cols_in + bk - 1) // bk, generator=g)
    i32 = lambda *s: torch.randint(-2 ** 31, 2 ** 31 - 1, s, generator=g, dtype=torch.int32)
    group = quant[1]
    if quant[0] == "awq":
        return {"qweight": lambda: i32(cols_in, rows_out // 8), "qzeros": lambda: i32(cols_in // group, rows_out // 8),
                "scales": lambda: torch.rand(cols_in // group, rows_out, generator=g).to(dtype)}[pname]()
    return {"qweight": lambda: i32(cols_in // 8, rows_out), "qzeros": lambda: i32(cols_in // group, rows_out // 8),
            "scales": lambda: torch.rand(cols_in // group, rows_out, generator=g).to(dtype),
            "g_idx": lambda: (torch.arange(cols_in, dtype=torch.int32) // group)}[pname]()


CASES = [
    # name, layer kind, layer args, quant, dtype, tp sizes
    ("merged_plain", "merged", dict(hidden=32, outs=[48, 48]), None, torch.float32, [1, 2, 4]),
    ("merged_uneven", "merged", dict(hidden=16, outs=[32, 64]), None, torch.bfloat16, [1, 2]),
    ("merged_awq", "merged", dict(hidden=256, outs=[64, 64]), ("awq", 128), torch.float16, [1, 2]),
    ("merged_fp8_tensor", "merged", dict(hidden=32, outs=[64, 64]), ("fp8_tensor",), torch.bfloat16, [1, 2]),
    ("merged_fp8_block", "merged", dict(hidden=256, outs=[512, 512]), ("fp8_block", [128, 128]), torch.bfloat16, [1, 2]),
    ("qkv_plain", "qkv", dict(hidden=24, d=4, heads=8, kv_heads=2), None, torch.float32, [1, 2, 4, 8]),
    ("qkv_awq", "qkv", dict(hidden=128, d=16, heads=8, kv_heads=2), ("awq", 64), torch.float16, [1, 2, 4]),
    ("qkv_fp8_tensor", "qkv", dict(hidden=32, d=4, heads=8, kv_heads=2), ("fp8_tensor",), torch.bfloat16, [2]),
    ("qkv_fp8_block", "qkv", dict(hidden=256, d=128, heads=8, kv_heads=2), ("fp8_block", [128, 128]), torch.bfloat16, [1, 2]),
    ("row_plain", "row", dict(in_size=64, out_size=24), None, torch.float32, [1, 2, 4]),
    ("row_awq", "row", dict(in_size=256, out_size=32), ("awq", 64), torch.float16, [1, 2]),
    ("row_gptq", "row", dict(in_size=256, out_size=32), ("gptq", 64), torch.float16, [1, 2]),
    ("row_fp8_tensor", "row", dict(in_size=64, out_size=32), ("fp8_tensor",), torch.bfloat16, [2]),
    ("column_plain", "column", dict(in_size=16, out_size=48), None, torch.float32, [1, 2, 4]),
]


def projections(kind, args):
    """-> [(shard id as the checkpoint loader passes it, total output rows)]"""
    if kind == "merged":
        return [(i, n) for i, n in enumerate(args["outs"])]
    if kind == "qkv":
        d = args["d"]
        return [("q", args["heads"] * d), ("k", args["kv_heads"] * d), ("v", args["kv_heads"] * d)]
    return [(None, args["out_size"])]


def run_case(P, classes, name, kind, args, quant, dtype, tp):
    g = torch.Generator().manual_seed(zlib.crc32(name.encode()) % 1000 + 7)
    in_size = args.get("hidden", args.get("in_size"))
    projs = projections(kind, args)
    pnames = list(make_params(P, quant, 8 * 128, 8 * 128, len(projs), dtype))
    if quant and quant[0] == "gptq" and kind != "row":
        raise ValueError("gptq case is row-parallel only here")
    ckpt = {(sid, pn): checkpoint_tensor(g, pn, quant, rows, in_size, dtype) for sid, rows in projs for pn in pnames}
    ranks = []
    for rank in range(tp):
        block = quant[1] if quant and quant[0] == "fp8_block" else None
        if kind == "merged":
            layer = ref_layer(classes["MergedColumnParallelLinear"], rank, tp, block, output_sizes=args["outs"])
            in_pp, out_pp = in_size, sum(args["outs"]) // tp
        elif kind == "qkv":
            f = qkv_fields(args["heads"], args["kv_heads"], args["d"], tp)
            layer = ref_layer(classes["QKVParallelLinear"], rank, tp, block, **f)
            in_pp, out_pp = in_size, (f["num_heads"] + 2 * f["num_kv_heads"]) * args["d"]
        elif kind == "row":
            layer = ref_layer(classes["RowParallelLinear"], rank, tp, block)
            in_pp, out_pp = in_size // tp, args["out_size"]
        else:
            layer = ref_layer(classes["ColumnParallelLinear"], rank, tp, block)
            in_pp, out_pp = in_size, args["out_size"] // tp
        params = make_params(P, quant, in_pp, out_pp, len(projs), dtype)
        for (sid, pn), t in ckpt.items():
            if sid is None:
                layer.weight_loader_v2(params[pn], t.clone())
            else:
                layer.weight_loader_v2(params[pn], t.clone(), sid)
        out = {pn: p.data.clone() for pn, p in params.items()}
        # the same projections fused on disk (no shard id): tensors with an output dim only
        fused = {}
        if kind in ("merged", "qkv"):
            fresh = make_params(P, quant, in_pp, out_pp, len(projs), dtype)
            for pn, p in fresh.items():
                # per-tensor scalars have no output dim; fused-on-disk block-scale grids are not
                # handled by the reference (linear.py:680-694 narrows with un-blocked sizes)
                if isinstance(p, (P.PerTensorScaleParameter, P.BlockQuantScaleParameter)):
                    continue
                cat = torch.cat([ckpt[(sid, pn)] for sid, _ in projs], dim=p.output_dim)
                layer.weight_loader_v2(p, cat)
                fused[pn] = p.data.clone()
                assert torch.equal(fused[pn].view(torch.uint8) if fused[pn].dtype == torch.float8_e4m3fn else fused[pn],
                                   out[pn].view(torch.uint8) if out[pn].dtype == torch.float8_e4m3fn else out[pn]), (name, pn)
        ranks.append({"expected": out, "fused_params": sorted(fused)})
    return {"name": name, "kind": kind, "args": args, "quant": quant, "dtype": str(dtype), "tp": tp,
            "checkpoint": {f"{sid}|{pn}": t for (sid, pn), t in ckpt.items()}, "ranks": ranks}


def main():
    P, classes = load_reference()
    out = []
    for name, kind, args, quant, dtype, tps in CASES:
        for tp in tps:
            c = run_case(P, classes, name, kind, args, quant, dtype, tp)
            out.append(c)
            print(name, "tp", tp, {k: tuple(v.shape) for k, v in c["ranks"][-1]["expected"].items()})
    # convert_to_channelwise (quantization/utils.py:68-88): what Fp8LinearMethod.process_weights_after_loading
    # does with the per-shard scales of an fp8-serialized checkpoint on the per-channel kernel path (fp8.py:364-370)
    uns = {"List": List, "Tuple": Tuple, "torch": torch}
    upath = f"{REF}/quantization/utils.py"
    fn = [n for n in ast.parse(open(upath).read()).body
          if isinstance(n, ast.FunctionDef) and n.name == "convert_to_channelwise"]
    exec(compile(ast.Module(fn, []), upath, "exec"), uns)
    channelwise = []
    for scales, widths in [([0.5, 0.25, 2.0], [32, 8, 8]), ([0.125, 3.0], [48, 48]), ([1.5], [64]), (0.75, [16, 4, 4])]:
        t = torch.tensor(scales, dtype=torch.float32)
        channelwise.append({"weight_scale": t, "logical_widths": widths,
                            "expected": uns["convert_to_channelwise"](t.clone(), widths)})
    path = os.path.join(ROOT, "tests", "golden", "weight_loaders.pt")
    torch.save({"cases": out, "channelwise": channelwise, "source": "reference parameter.py + linear.py weight_loader_v2, torch " + torch.__version__}, path)
    print(path, os.path.getsize(path))


if __name__ == "__main__":
    sys.exit(main())
