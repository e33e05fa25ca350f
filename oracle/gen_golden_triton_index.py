"""Generates tests/golden/triton_index.pt by running two more of the reference's OWN Triton kernels
through Triton's CPU interpreter (TRITON_INTERPRET=1), source text extracted unmodified:
  create_flashinfer_kv_indices_triton  python/sglang/srt/layers/attention/utils.py:5-41
  write_req_to_token_pool_triton       python/sglang/srt/managers/schedule_batch.py:1789-1823
launched as the reference launches them (triton_backend.py:200-240 / test_create_kvindices.py:52-60;
schedule_batch.py:1272-1283).  Run here (needs /root/reference and triton).  TEST INFRASTRUCTURE ONLY."""
import ast
import importlib.util
import os
import sys
import tempfile

os.environ["TRITON_INTERPRET"] = "1"   # before triton is imported

import torch

REF = "/root/reference/python/sglang/srt"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def load_kernel(path, name):
    src = open(path).read()
    lines = src.splitlines()
    for node in ast.parse(src).body:
        if isinstance(node, ast.FunctionDef) and node.name == name:
            start = min([d.lineno for d in node.decorator_list] + [node.lineno]) - 1
            text = "\n".join(lines[start:node.end_lineno])
            mod_path = os.path.join(tempfile.mkdtemp(), f"ref_{name}.py")
            open(mod_path, "w").write("import triton\nimport triton.language as tl\n\n\n" + text + "\n")
            spec = importlib.util.spec_from_file_location(f"ref_{name}", mod_path)
            mod = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(mod)
            return getattr(mod, name)
    raise KeyError(name)


def main():
    kv_indices_kernel = load_kernel(f"{REF}/layers/attention/utils.py", "create_flashinfer_kv_indices_triton")
    write_kernel = load_kernel(f"{REF}/managers/schedule_batch.py", "write_req_to_token_pool_triton")
    g = torch.Generator().manual_seed(0)
    kv_cases, write_cases = [], []
    for bs, ctx, with_start in [(1, 600, False), (7, 1500, False), (37, 520, True), (64, 64, True)]:
        max_batch = bs + 5
        rtt = torch.randint(1, 1 << 20, (max_batch, ctx), generator=g, dtype=torch.int32)
        rpi = torch.randperm(max_batch, generator=g)[:bs].to(torch.int32)
        start = (torch.randint(0, ctx // 3, (bs,), generator=g, dtype=torch.int32) if with_start else None)
        room = ctx - (start if with_start else torch.zeros(bs, dtype=torch.int32))
        lens = (torch.rand(bs, generator=g) * room).to(torch.int32)
        lens[0] = room[0]                       # one request uses its whole row
        if bs > 2:
            lens[1] = 0                         # an empty request
        indptr = torch.zeros(bs + 1, dtype=torch.int32)
        indptr[1:] = torch.cumsum(lens, 0)
        out = torch.full((int(indptr[-1]),), -1, dtype=torch.int32)
        kv_indices_kernel[(bs,)](rtt, rpi, lens, indptr, start, out, rtt.stride(0))
        kv_cases.append(dict(req_to_token=rtt, req_pool_indices=rpi, lens=lens, kv_start=start,
                             kv_indptr=indptr, kv_indices=out))
    for bs, ctx in [(1, 700), (9, 1300), (64, 600)]:
        max_batch = bs + 3
        rtt = torch.randint(1, 1 << 20, (max_batch, ctx), generator=g, dtype=torch.int32)
        before = rtt.clone()
        rpi = torch.randperm(max_batch, generator=g)[:bs].to(torch.int64)
        pre = torch.randint(0, ctx // 2, (bs,), generator=g, dtype=torch.int64)
        ext = torch.randint(1, ctx // 2, (bs,), generator=g, dtype=torch.int64)
        pre[0] = 0
        seq = pre + ext
        loc = torch.randint(1, 1 << 22, (int(ext.sum()),), generator=g, dtype=torch.int64)
        write_kernel[(bs,)](rtt, rpi, pre, seq, ext, loc, rtt.stride(0))
        write_cases.append(dict(req_to_token_before=before, req_pool_indices=rpi, pre_lens=pre, seq_lens=seq,
                                extend_lens=ext, out_cache_loc=loc, req_to_token_after=rtt))
    out_path = os.path.join(ROOT, "tests", "golden", "triton_index.pt")
    torch.save(dict(kv_indices=kv_cases, write_req_to_token=write_cases), out_path)
    print(out_path, len(kv_cases), len(write_cases))


if __name__ == "__main__":
    sys.exit(main())
