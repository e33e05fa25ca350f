"""Generates tests/golden/sched_alloc.pt by running the reference's OWN allocator kernels:
`alloc_extend_kernel` / `alloc_decode_kernel` of python/sglang/srt/mem_cache/paged_allocator.py:29-155.
They are Triton kernels; their source text is extracted UNMODIFIED (ast line ranges, decorators
included) into a scratch module and executed on CPU tensors with Triton's interpreter
(TRITON_INTERPRET=1), launched exactly as PagedTokenToKVPoolAllocator.alloc_extend / alloc_decode
launch them (paged_allocator.py:217-262).  Run here (needs /root/reference and triton); the fixture
travels to the GPU box.  TEST INFRASTRUCTURE ONLY."""
import ast
import importlib.util
import os
import sys
import tempfile

os.environ["TRITON_INTERPRET"] = "1"   # before triton is imported

import torch

REF = "/root/reference/python/sglang/srt/mem_cache/paged_allocator.py"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def load_kernels():
    src = open(REF).read()
    lines = src.splitlines()
    segs = []
    for node in ast.parse(src).body:
        if isinstance(node, ast.FunctionDef) and node.name in ("alloc_extend_kernel", "alloc_decode_kernel"):
            start = min([d.lineno for d in node.decorator_list] + [node.lineno]) - 1
            segs.append("\n".join(lines[start:node.end_lineno]))
    assert len(segs) == 2
    path = os.path.join(tempfile.mkdtemp(), "ref_alloc_kernels.py")
    open(path, "w").write("import triton\nimport triton.language as tl\n\n\n" + "\n\n\n".join(segs) + "\n")
    spec = importlib.util.spec_from_file_location("ref_alloc_kernels", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.alloc_extend_kernel, mod.alloc_decode_kernel


def next_power_of_2(n):  # sglang.srt.utils.next_power_of_2
    return 1 << (n - 1).bit_length() if n > 0 else 1


def main():
    alloc_extend_kernel, alloc_decode_kernel = load_kernels()
    g = torch.Generator().manual_seed(0)
    cases = []
    for page, bs, max_pre, max_ext in [(1, 5, 40, 30), (16, 1, 100, 50), (16, 9, 200, 90), (64, 33, 300, 200),
                                       (16, 64, 64, 1), (32, 7, 0, 100)]:
        pre = torch.randint(0, max_pre + 1, (bs,), generator=g, dtype=torch.int64)
        ext = torch.randint(1, max_ext + 1, (bs,), generator=g, dtype=torch.int64)
        if bs > 2:
            pre[0] = 0                       # a request without prefix
            pre[1] = page * 3                # a page-aligned prefix
            ext[2] = (page - pre[2] % page) % page or page   # ends exactly on a page boundary
        seq = pre + ext
        # last_loc: slot of the last prefix token; must sit at offset (pre - 1) % page of some page
        # (the invariant the reference asserts in debug mode), -1 for an empty prefix
        owned = torch.randperm(4096, generator=g)[:bs].to(torch.int64) + 8192
        last = torch.where(pre > 0, owned * page + (pre - 1) % page, torch.full_like(pre, -1))
        free_pages = torch.randperm(4000, generator=g).to(torch.int64) + 1
        n_ext = int(ext.sum())
        out = torch.empty((n_ext,), dtype=torch.int64)
        ret = torch.empty((), dtype=torch.int64)
        alloc_extend_kernel[(bs,)](pre, seq, last, free_pages, out, ret, next_power_of_2(bs), page,
                                   next_power_of_2(n_ext))
        merged = int(ret.item())
        # decode on the state after the extend: every request grows by one token
        seq_d = seq + 1
        # last token of each request as the extend just placed it
        start = torch.cumsum(ext, 0) - ext
        last_d = out[start + ext - 1]
        free_after = free_pages[merged >> 32:]
        out_d = torch.empty((bs,), dtype=torch.int64)
        ret_d = torch.empty((), dtype=torch.int64)
        alloc_decode_kernel[(bs,)](seq_d, last_d, free_after, out_d, ret_d, next_power_of_2(bs), page)
        cases.append(dict(page_size=page, prefix_lens=pre, seq_lens=seq, last_loc=last, free_pages=free_pages,
                          extend_out_indices=out, extend_num_new_pages=merged >> 32,
                          extend_sum_extend_lens=merged & 0xFFFFFFFF,
                          decode_seq_lens=seq_d, decode_last_loc=last_d, decode_free_pages=free_after,
                          decode_out_indices=out_d, decode_num_new_pages=int(ret_d.item())))
    out_path = os.path.join(ROOT, "tests", "golden", "sched_alloc.pt")
    torch.save(cases, out_path)
    print(out_path, len(cases))


if __name__ == "__main__":
    sys.exit(main())
