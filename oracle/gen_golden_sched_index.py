"""Generates tests/golden/sched_index.pt by running the reference's OWN torch implementations,
whose source is extracted unmodified (ast) from the read-only reference tree:
  get_last_loc_torch      python/sglang/srt/managers/schedule_batch.py
  compute_position_torch  python/sglang/srt/model_executor/forward_batch_info.py
Run here (needs /root/reference); the fixture travels to the GPU box.  TEST INFRASTRUCTURE ONLY."""
import ast
import os
import sys

import torch

REF = "/root/reference/python/sglang/srt"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def load_func(path, name):
    src = open(path).read()
    for node in ast.parse(src).body:
        if isinstance(node, ast.FunctionDef) and node.name == name:
            node.decorator_list = []
            ns = {"torch": torch}
            exec(compile(ast.Module([node], []), path, "exec"), ns)
            return ns[name]
    raise KeyError(name)


def main():
    get_last_loc_torch = load_func(f"{REF}/managers/schedule_batch.py", "get_last_loc_torch")
    compute_position_torch = load_func(f"{REF}/model_executor/forward_batch_info.py",
                                       "compute_position_torch")
    g = torch.Generator().manual_seed(0)
    cases = []
    for bs, ctx in [(1, 16), (7, 300), (64, 520), (300, 64)]:
        rtt = torch.randint(1, 1 << 20, (bs + 3, ctx), generator=g, dtype=torch.int32)
        rpi = torch.randperm(bs + 3, generator=g)[:bs].to(torch.int64)
        prefix = torch.randint(0, ctx, (bs,), generator=g, dtype=torch.int64)
        prefix[0] = 0
        ext = torch.randint(1, 40, (bs,), generator=g, dtype=torch.int32)
        last = get_last_loc_torch(rtt, rpi, prefix)
        pos, start = compute_position_torch(prefix.to(torch.int32), ext)
        cases.append(dict(req_to_token=rtt, req_pool_indices=rpi, prefix_lens=prefix,
                          extend_seq_lens=ext, last_loc=last, positions=pos,
                          extend_start_loc=start))
    out = os.path.join(ROOT, "tests", "golden", "sched_index.pt")
    torch.save(cases, out)
    print(out, len(cases))


if __name__ == "__main__":
    sys.exit(main())
