"""Generates tests/golden/kv_write_fp8.pt by running the reference's OWN
MHATokenToKVPool.set_kv_buffer / get_key_buffer / get_value_buffer on CPU.  The three methods are
extracted unmodified (ast) from the read-only reference tree
(python/sglang/srt/mem_cache/memory_pool.py:359-412) and bound to a bare object carrying the fields
the constructor would set (memory_pool.py:119-123,264-300); the one import inside set_kv_buffer
(cuda_graph_runner.get_is_capture_mode) is answered by a stub module returning False, which is its
value outside graph capture.  Covers float8_e4m3fn and float8_e5m2 pools, bf16 / f16 activations,
with and without k_scale / v_scale, out-of-range values (e4m3: NaN byte; e5m2: Inf byte), subnormals.
Run here (needs /root/reference); the fixture travels to the GPU box.  TEST INFRASTRUCTURE ONLY."""
import ast
import os
import sys
import types
from types import SimpleNamespace
from typing import Optional

import torch

REF = "/root/reference/python/sglang/srt/mem_cache/memory_pool.py"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def load_methods(path, cls, names):
    for node in ast.parse(open(path).read()).body:
        if isinstance(node, ast.ClassDef) and node.name == cls:
            fns = [n for n in node.body if isinstance(n, ast.FunctionDef) and n.name in names]
            for f in fns:
                f.decorator_list = []
                for a in f.args.args:
                    a.annotation = None      # RadixAttention annotation: not importable, not needed
            ns = {"torch": torch, "Optional": Optional}
            exec(compile(ast.Module(fns, []), path, "exec"), ns)
            return {n: ns[n] for n in names}
    raise KeyError(cls)


def reference_pool(methods, size, dtype, hkv, d):
    stub = types.ModuleType("sglang.srt.model_executor.cuda_graph_runner")
    stub.get_is_capture_mode = lambda: False
    for name in ("sglang", "sglang.srt", "sglang.srt.model_executor"):
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.modules["sglang.srt.model_executor.cuda_graph_runner"] = stub
    store = torch.uint8 if dtype in (torch.float8_e5m2, torch.float8_e4m3fn) else dtype
    pool = SimpleNamespace(dtype=dtype, store_dtype=store, start_layer=0, alt_stream=None,
                           layer_transfer_counter=None,
                           k_buffer=[torch.zeros((size + 1, hkv, d), dtype=store)],
                           v_buffer=[torch.zeros((size + 1, hkv, d), dtype=store)])
    for n, f in methods.items():
        setattr(pool, n, types.MethodType(f, pool))
    return pool


def make_inputs(seed, dtype, pool_dtype, t=29, hkv=2, d=64, size=100):
    g = torch.Generator().manual_seed(seed)
    k = torch.randn((t, hkv, d), generator=g) * 6
    v = torch.randn((t, hkv, d), generator=g) * (0.01 if pool_dtype == torch.float8_e4m3fn else 4e-5)
    if pool_dtype == torch.float8_e4m3fn:
        edge = [448., 449., 464., 465., 480., -1000., 0., -0.]
    else:
        edge = [57344., 59000., 61184., 61440., 63488., 65504., -1e5, float("inf"), float("-inf"),
                0., -0., 7.6e-6, 7.7e-6, -2.3e-5]
    k[0, 0, :len(edge)] = torch.tensor(edge)
    k[1, 0, :len(edge)] = torch.tensor(edge) * 0.5
    loc = (torch.randperm(size, generator=g)[:t] + 1).to(torch.int64)
    return k.to(dtype), v.to(dtype), loc


def main():
    methods = load_methods(REF, "MHATokenToKVPool", ["set_kv_buffer", "get_key_buffer", "get_value_buffer"])
    cases = []
    seed = 0
    for pool_dtype in (torch.float8_e4m3fn, torch.float8_e5m2):
        for dtype in (torch.bfloat16, torch.float16):
            for scales in (None, (0.5, 2.0), (0.0373, 1.7)):
                seed += 1
                k, v, loc = make_inputs(seed, dtype, pool_dtype)
                pool = reference_pool(methods, 100, pool_dtype, 2, 64)
                ks, vs = scales if scales else (None, None)
                pool.set_kv_buffer(SimpleNamespace(layer_id=0), loc, k.clone(), v.clone(), ks, vs)
                assert pool.get_key_buffer(0).dtype == pool_dtype
                cases.append(dict(pool_dtype=str(pool_dtype), dtype=str(dtype), scales=scales, k=k, v=v,
                                  loc=loc, k_rows=pool.k_buffer[0][loc].clone(),
                                  v_rows=pool.v_buffer[0][loc].clone()))
                untouched = torch.ones(101, dtype=torch.bool)
                untouched[loc] = False
                assert not pool.k_buffer[0][untouched].any() and not pool.v_buffer[0][untouched].any()
    out = os.path.join(ROOT, "tests", "golden", "kv_write_fp8.pt")
    torch.save({"cases": cases, "source": "reference memory_pool.py MHATokenToKVPool.set_kv_buffer on CPU, torch "
                + torch.__version__}, out)
    print(out, len(cases), os.path.getsize(out))


if __name__ == "__main__":
    sys.exit(main())
