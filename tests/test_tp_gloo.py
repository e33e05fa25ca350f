"""CPU, world_size 2 over gloo: the tensor-parallel host logic of the linear layers — column
sharding of N, row sharding of K, bias on rank 0 only, all-reduce at the row-parallel boundary
(reference linear.py:352-358, 1184, 1277-1299) — and head sharding of the attention path."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from sglang_security_b200.distributed import parallel_state as ps
    from sglang_security_b200.layers.linear import (MergedColumnParallelLinear, QKVParallelLinear,
                                                    RowParallelLinear)

    ps.initialize_model_parallel(world)
    assert ps.get_tensor_model_parallel_world_size() == world
    assert ps.get_tensor_model_parallel_rank() == rank
    torch.manual_seed(0)  # same full weights on every rank
    hidden, inter, heads, kv_heads, d = 64, 96, 4, 2, 16
    x = torch.randn(5, hidden)
    w_gate_up = torch.randn(2 * inter, hidden)
    w_down = torch.randn(hidden, inter)
    b_down = torch.randn(hidden)
    w_qkv = torch.randn((heads + 2 * kv_heads) * d, hidden)

    gate_up = MergedColumnParallelLinear(hidden, [inter, inter], bias=False, params_dtype=torch.float32)
    down = RowParallelLinear(inter, hidden, bias=True, params_dtype=torch.float32)
    qkv = QKVParallelLinear(hidden, d, heads, kv_heads, bias=False, params_dtype=torch.float32)
    assert gate_up.weight.shape == (2 * inter // world, hidden)
    assert down.weight.shape == (hidden, inter // world)
    assert qkv.num_heads == heads // world and qkv.num_kv_heads == kv_heads // world
    # shard-wise loading of fused weights: gate rows then up rows of THIS rank
    shard = inter // world
    gate_up.weight.data[:shard].copy_(w_gate_up[:inter][rank * shard:(rank + 1) * shard])
    gate_up.weight.data[shard:].copy_(w_gate_up[inter:][rank * shard:(rank + 1) * shard])
    down.weight_loader(down.weight, w_down)
    down.bias.data.copy_(b_down)

    h, _ = gate_up(x)
    act = torch.nn.functional.silu(h[:, :shard]) * h[:, shard:]
    y, _ = down(act)                      # all-reduce inside
    full_h = x @ w_gate_up.t()
    ref = (torch.nn.functional.silu(full_h[:, :inter]) * full_h[:, inter:]) @ w_down.t() + b_down
    torch.testing.assert_close(y, ref, rtol=1e-4, atol=1e-4)

    # all-gather of a column-parallel output (logits path, linear.py:446-448)
    g = ps.tensor_model_parallel_all_gather(h[:, :shard].contiguous())
    torch.testing.assert_close(g, full_h[:, :inter], rtol=1e-4, atol=1e-4)
    dist.barrier()
    open(os.path.join(out_dir, f"ok{rank}"), "w").write("ok")
    dist.destroy_process_group()


def test_tp2_linear_layers_gloo(tmp_path):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    assert all((tmp_path / f"ok{r}").exists() for r in range(world))


def test_bench_reference_arm_nonzero_ranks_exit_quietly(monkeypatch, capsys):
    """`bench.py --impl reference` under torchrun: only rank 0 works and prints."""
    import bench

    monkeypatch.setenv("RANK", "1")
    bench.run_reference(type("A", (), {"steps": 1, "warmup": 1, "gpus": 2})())
    assert capsys.readouterr().out == ""
