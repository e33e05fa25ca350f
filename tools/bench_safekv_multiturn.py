#!/usr/bin/env python
"""BASELINE cfg 5: multi-turn shared-prefix load with the SafeKV private / shared prefix split,
radix hit-rate sweep — prefill (extend) and decode tokens/s of the whole model against hit rate.

The SafeKV policy itself stays on the host, untouched (north_star): this tool only MODELS its
reuse rule to produce the (extend_prefix_lens, extend_seq_lens) the scheduler would hand the model
runner, and then runs the real B200 hot path on them.  Rule restated from the fork's
RadixCache._match_prefix_helper (python/sglang/srt/mem_cache/radix_cache.py:333-349): a cached node
is reusable by everyone unless it is `private`, in which case only the request whose
`user_id == node.owner_id` may match through it; matching stops at the first refused node.

Workload: U users x T turns.  Every prompt = shared system prompt (S tokens, public node) + the
user's own history (private nodes, one per past turn: question Q + answer A tokens) + a new
question.  With probability `--foreign` a turn is re-issued under ANOTHER user id (same tokens — a
colleague opening the same conversation): SafeKV then refuses the private history and only the
system prompt hits.  foreign = 0 / 0.5 / 1 sweeps the hit rate from high to low.

Per turn-batch (all users at turn t): one EXTEND step over the cached prefixes (tcgen05 extend
attention + large-M FP8 GEMMs), then `--decode-steps` DECODE steps at the resulting context.
Run under torchrun for TP > 1 (Llama-3.1-70B FP8 at TP 8 is the cfg-5 point); on one GPU the
default is Llama-3.1-8B.  Prints one JSON line per foreign-probability point.
"""
import argparse
import json
import os
import random
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch
import torch.distributed as dist


class SafeKVPrefixModel:
    """Block-granular stand-in for the fork's radix tree: a path of (block id) nodes per prompt."""

    def __init__(self):
        self.nodes = {}  # path tuple -> (private, owner_id)

    def match(self, blocks, user_id):
        """-> number of leading blocks reusable by `user_id` (radix_cache.py:342-344)."""
        n = 0
        for i in range(len(blocks)):
            node = self.nodes.get(tuple(blocks[: i + 1]))
            if node is None:
                break
            private, owner = node
            if private and owner != user_id:
                break
            n += 1
        return n

    def insert(self, blocks, user_id, private_from):
        for i in range(len(blocks)):
            key = tuple(blocks[: i + 1])
            if key not in self.nodes:
                self.nodes[key] = (i >= private_from, user_id)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="")
    ap.add_argument("--users", type=int, default=32)
    ap.add_argument("--turns", type=int, default=4)
    ap.add_argument("--system", type=int, default=1024)
    ap.add_argument("--question", type=int, default=256)
    ap.add_argument("--answer", type=int, default=128)
    ap.add_argument("--decode-steps", type=int, default=8, help="timed decode steps per turn batch")
    ap.add_argument("--foreign", default="0,0.5,1")
    ap.add_argument("--layers", type=int, default=0)
    ap.add_argument("--out", default="")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    from sglang_security_b200.distributed.parallel_state import (destroy_model_parallel,
                                                                initialize_model_parallel)
    from sglang_security_b200.model_executor.decode_step import GEOMETRIES, B200DecodeStep
    from sglang_security_b200.model_executor.forward_batch_info import ForwardBatch, ForwardMode

    initialize_model_parallel(world)
    name = args.model or ("llama3.1-70b" if world >= 8 else "llama3.1-8b")
    geom = GEOMETRIES[name]
    U, T, S, Q, A = args.users, args.turns, args.system, args.question, args.answer
    max_ctx = S + T * (Q + A) + 8
    model = B200DecodeStep(geom, "fp8", dev, tp_rank=rank, tp_size=world, max_bs=U, max_ctx=max_ctx,
                           pool_tokens=U * max_ctx, layers=args.layers, fill_kv=False)
    rtt = model.req_to_token_pool.req_to_token
    # request u owns the slot range [1 + u * max_ctx, ...): position p of request u lives in slot base_u + p
    base = torch.arange(U, dtype=torch.int64) * max_ctx + 1
    rtt[:U, :max_ctx] = (base.view(U, 1) + torch.arange(max_ctx).view(1, max_ctx)).to(torch.int32).to(dev)
    vocab = geom.vocab
    rows = []
    for foreign in [float(x) for x in args.foreign.split(",")]:
        rng = random.Random(1234)
        tree = SafeKVPrefixModel()
        g = torch.Generator().manual_seed(7)
        cached = prompt = new_tokens = 0
        t_prefill = t_decode = 0.0
        dec_tokens = 0
        for turn in range(T):
            ctx_before = S + turn * (Q + A)              # tokens of every user's history so far
            prefix_lens, extend_lens = [], []
            for u in range(U):
                blocks = ["sys"] + [f"u{u}t{i}" for i in range(turn)]
                sizes = [S] + [Q + A] * turn
                uid = u if rng.random() >= foreign else (u + 1) % U + U   # a foreign id never owns anything
                hit_blocks = tree.match(blocks, uid)
                hit = sum(sizes[:hit_blocks])
                prefix_lens.append(hit)
                extend_lens.append(ctx_before + Q - hit)
                tree.insert(blocks, u, private_from=1)   # history is private to its owner, "sys" is shared
            cached += sum(prefix_lens)
            prompt += U * (ctx_before + Q)
            n_new = sum(extend_lens)
            new_tokens += n_new
            # ---- extend step ------------------------------------------------------------------
            pos = torch.cat([torch.arange(p, p + e) for p, e in zip(prefix_lens, extend_lens)])
            req = torch.cat([torch.full((e,), u, dtype=torch.int64) for u, e in enumerate(extend_lens)])
            loc = (base[req] + pos).to(dev)
            ids = torch.randint(0, vocab, (n_new,), generator=g).to(dev)
            ext = torch.tensor(extend_lens, dtype=torch.int32)
            pre = torch.tensor(prefix_lens, dtype=torch.int32)
            start = torch.zeros(U, dtype=torch.int32)
            start[1:] = torch.cumsum(ext[:-1], 0)
            seq = (pre + ext).to(torch.int64)
            fb = ForwardBatch(forward_mode=ForwardMode.EXTEND, batch_size=U,
                              req_pool_indices=torch.arange(U, dtype=torch.int64, device=dev),
                              seq_lens=seq.to(dev), out_cache_loc=loc, seq_lens_sum=int(seq.sum()),
                              seq_lens_cpu=seq, extend_seq_lens=ext.to(dev), extend_prefix_lens=pre.to(dev),
                              extend_start_loc=start.to(dev), extend_seq_lens_cpu=extend_lens,
                              extend_prefix_lens_cpu=prefix_lens, extend_num_tokens=n_new,
                              req_to_token_pool=model.req_to_token_pool,
                              token_to_kv_pool=model.token_to_kv_pool, attn_backend=model.attn_backend)
            last = (start + ext - 1).to(torch.int64).to(dev)
            out = torch.empty((U,), dtype=torch.int64, device=dev)
            model.attn_backend.init_forward_metadata(fb)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            if turn == 0 and foreign == float(args.foreign.split(",")[0]):   # one untimed warm-up
                model.forward_extend(ids, pos.to(dev), fb, last, out)
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()
            e0.record()
            model.forward_extend(ids, pos.to(dev), fb, last, out)
            e1.record()
            torch.cuda.synchronize()
            t_prefill += e0.elapsed_time(e1)
            # ---- decode steps at the new context --------------------------------------------
            ctx = ctx_before + Q
            for step in range(args.decode_steps):
                p = ctx + step
                d_pos = torch.full((U,), p, dtype=torch.int64, device=dev)
                d_loc = (base + p).to(dev)
                d_len = torch.full((U,), p + 1, dtype=torch.int64, device=dev)
                fbd = model.make_forward_batch(U, torch.arange(U, dtype=torch.int64, device=dev), d_len,
                                               d_loc, U * (p + 1))
                model.attn_backend.init_forward_metadata(fbd)
                e0.record()
                model.forward(out, d_pos, fbd, out)
                e1.record()
                torch.cuda.synchronize()
                if step > 0:                      # the first step of a turn pays allocator warm-up
                    t_decode += e0.elapsed_time(e1)
                    dec_tokens += U
        times = torch.tensor([t_prefill, t_decode], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(times, op=dist.ReduceOp.MAX)
        t_prefill, t_decode = float(times[0]), float(times[1])
        row = {"model": name, "tp": world, "users": U, "turns": T, "system": S, "question": Q, "answer": A,
               "foreign_user_prob": foreign, "radix_hit_rate": round(cached / prompt, 4),
               "prompt_tokens": prompt, "extend_tokens": new_tokens,
               "prefill_ms": round(t_prefill, 2),
               "prefill_new_tokens_per_s": round(new_tokens / (t_prefill * 1e-3), 1),
               "prefill_prompt_tokens_per_s": round(prompt / (t_prefill * 1e-3), 1),
               "decode_tokens_per_s": round(dec_tokens / (t_decode * 1e-3), 1) if t_decode else None,
               "decode_mode": "eager (per-step launches, no CUDA graph)", "layers": model.layers_n}
        rows.append(row)
        if rank == 0:
            print(json.dumps(row), flush=True)
    if rank == 0 and args.out:
        os.makedirs(os.path.dirname(os.path.abspath(args.out)), exist_ok=True)
        json.dump(rows, open(args.out, "w"), indent=1)
    torch.cuda.synchronize()
    if world > 1:
        destroy_model_parallel()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
