"""GPU parity of the NVLink peer-memory all-reduce and of the row-parallel GEMM that reduces in its
epilogue (csrc/comm.cu, csrc/fp8_gemm_smallm.cuh) against oracle/allreduce.py (rank-ordered fp32
sum, one rounding == the reference's one-shot kernel arithmetic) — bit-exact.

Single-GPU box: the `world` ranks are driven from `world` CUDA streams of one device
(LocalCommGroup: same kernels, buffers not IPC-mapped).  With >= 2 GPUs the multi-process tests run
the real cudaIpc path next to NCCL."""
import os
import socket

import pytest
import torch

from oracle import allreduce as OA

pytestmark = pytest.mark.gpu
FP8 = torch.float8_e4m3fn


def _group(world, max_size=4 << 20):
    from sglang_security_b200.distributed.device_communicators.b200_all_reduce import LocalCommGroup

    return LocalCommGroup(world, max_size=max_size)


def _run_ranks(grp, fn):
    """fn(rank, comm) on every rank's stream; returns the per-rank results after a device sync."""
    outs = []
    for r in range(grp.world):
        grp.streams[r].wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(grp.streams[r]):
            outs.append(fn(r, grp.comms[r]))
    torch.cuda.synchronize()
    return outs


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
def test_all_reduce_bit_exact_local(world, dtype):
    grp = _group(world)
    g = torch.Generator().manual_seed(world)
    try:
        # 16 B .. 768 KiB, several calls back to back (epoch / parity double-buffering)
        for numel in [8, 1032, 4096 * 8, 48 * 4096]:  # <= 96 CTAs per rank: all ranks co-resident
            xs = [(torch.randn(numel, generator=g) * 3).to(dtype) for _ in range(world)]
            ref = OA.all_reduce_oracle(xs)
            dev = [x.cuda() for x in xs]
            outs = _run_ranks(grp, lambda r, c: c.all_reduce(dev[r]))
            for r in range(world):
                assert torch.equal(outs[r].cpu(), ref), (numel, r)
            # in place
            outs = _run_ranks(grp, lambda r, c: c.all_reduce(dev[r], out=dev[r]))
            for r in range(world):
                assert torch.equal(outs[r].cpu(), ref), ("inplace", numel, r)
    finally:
        grp.close()


def test_all_reduce_rejects_bad_sizes():
    from sglang_security_b200 import _lib

    grp = _group(2, max_size=1 << 16)
    try:
        x = torch.zeros(3, dtype=torch.bfloat16, device="cuda")  # 6 bytes: not a multiple of 16
        with pytest.raises(_lib.B200Error):
            grp.comms[0].all_reduce(x)
        big = torch.zeros((1 << 16) + 8, dtype=torch.bfloat16, device="cuda")
        with pytest.raises(_lib.B200Error):
            grp.comms[0].all_reduce(big)
    finally:
        grp.close()


def test_all_reduce_cuda_graph_replay_local():
    """Epochs live in device memory: a captured all-reduce replays without host bookkeeping."""
    world = 2
    grp = _group(world)
    try:
        g = torch.Generator().manual_seed(5)
        dev = [torch.zeros(4096, dtype=torch.bfloat16, device="cuda") for _ in range(world)]
        outs = [torch.empty_like(d) for d in dev]
        _run_ranks(grp, lambda r, c: c.all_reduce(dev[r], out=outs[r]))  # warm-up
        graphs = []
        for r in range(world):
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr, stream=grp.streams[r]):
                grp.comms[r].all_reduce(dev[r], out=outs[r])
            graphs.append(gr)
        for it in range(3):
            xs = [torch.randn(4096, generator=g).to(torch.bfloat16) for _ in range(world)]
            for r in range(world):
                dev[r].copy_(xs[r])
            torch.cuda.synchronize()
            for r in range(world):
                with torch.cuda.stream(grp.streams[r]):
                    graphs[r].replay()
            torch.cuda.synchronize()
            ref = OA.all_reduce_oracle(xs)
            for r in range(world):
                assert torch.equal(outs[r].cpu(), ref), (it, r)
    finally:
        grp.close()


def _rand_fp8(shape, g, amp=448.0):
    return ((torch.rand(shape, generator=g) - 0.5) * 2 * amp).clamp(-448, 448).to(FP8)


@pytest.mark.parametrize("world,n,k", [(2, 4096, 2048), (2, 1024, 7168), (4, 1024, 1024),
                                       (8, 512, 512)])
@pytest.mark.parametrize("m", [1, 7, 16, 33, 64, 100, 128])
def test_fp8_rowcol_gemm_fused_all_reduce_local(world, n, k, m):
    """Row-parallel FP8 GEMM: fused epilogue all-reduce == oracle sum of the per-rank outputs.
    (M > 128 = big-tile GEMM followed by the stand-alone all-reduce is covered by the multi-process
    test only: with every rank on ONE device the early-launched, spinning all-reduce CTAs of rank 0
    can keep rank 1's GEMM from being scheduled.)"""
    from sglang_security_b200.layers.quantization.fp8_kernel import (comm_fuse_config,
                                                                    fp8_gemm_smallm_config,
                                                                    fp8_scaled_mm)

    grp = _group(world)
    # the fused path always takes the swap-AB kernel for M <= 128; make the un-fused per-rank
    # outputs come from the same kernel (same split-K grouping => same fp32 partial sums)
    fp8_gemm_smallm_config(max_m=128)
    comm_fuse_config(128)  # fuse for every M the kernel supports, whatever the tuned policy is
    try:
        g = torch.Generator().manual_seed(m * 31 + n + world)
        a = [_rand_fp8((m, k), g, 8.0).cuda() for _ in range(world)]
        w = [_rand_fp8((n, k), g, 4.0).cuda() for _ in range(world)]
        sa = [(torch.rand(m, generator=g) * 1e-2 + 1e-3).cuda() for _ in range(world)]
        sb = [(torch.rand(n, generator=g) * 1e-2 + 1e-3).cuda() for _ in range(world)]
        bias = (torch.randn(n, generator=g)).to(torch.bfloat16).cuda()
        # per-rank outputs of the un-fused kernel (itself pinned to oracle/fp8.py in test_fp8_gpu)
        parts = [fp8_scaled_mm(a[r], w[r].t(), sa[r], sb[r], torch.bfloat16,
                               bias=bias if r == 0 else None) for r in range(world)]
        torch.cuda.synchronize()
        ref = OA.row_parallel_linear_oracle(parts)
        for _ in range(2):  # twice: both parities of the receive slots
            outs = _run_ranks(grp, lambda r, c: fp8_scaled_mm(
                a[r], w[r].t(), sa[r], sb[r], torch.bfloat16, bias=bias if r == 0 else None, comm=c))
            for r in range(world):
                assert torch.equal(outs[r].cpu(), ref), (r, (outs[r].cpu().float() - ref.float()).abs().max())
    finally:
        fp8_gemm_smallm_config(-1, -1)
        comm_fuse_config(-1)
        grp.close()


@pytest.mark.parametrize("world,n,k", [(2, 4096, 1024), (4, 1024, 2048)])
@pytest.mark.parametrize("m", [1, 16, 48, 64])
def test_fp8_block_gemm_fused_all_reduce_local(world, n, k, m):
    from sglang_security_b200.layers.quantization.fp8_kernel import (comm_fuse_config,
                                                                    fp8_blockwise_scaled_mm)

    grp = _group(world)
    comm_fuse_config(128)
    try:
        g = torch.Generator().manual_seed(m * 17 + n + world)
        a = [_rand_fp8((m, k), g, 8.0).cuda() for _ in range(world)]
        w = [_rand_fp8((n, k), g, 4.0).cuda() for _ in range(world)]
        as_ = [(torch.rand((m, k // 128), generator=g) * 1e-2 + 1e-3).cuda() for _ in range(world)]
        bs = [(torch.rand((n // 128, k // 128), generator=g) * 1e-2 + 1e-3).cuda()
              for _ in range(world)]
        parts = [fp8_blockwise_scaled_mm(a[r], w[r], as_[r], bs[r], torch.bfloat16)
                 for r in range(world)]
        torch.cuda.synchronize()
        ref = OA.row_parallel_linear_oracle(parts)
        outs = _run_ranks(grp, lambda r, c: fp8_blockwise_scaled_mm(
            a[r], w[r], as_[r], bs[r], torch.bfloat16, comm=c))
        for r in range(world):
            assert torch.equal(outs[r].cpu(), ref), r
    finally:
        comm_fuse_config(-1)
        grp.close()


# ---- real multi-process path (cudaIpc + NCCL side by side): needs >= 2 GPUs ---------------------
def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _mp_worker(rank, world, port, out_dir, nvls=False):
    import torch.distributed as dist

    os.environ["B200_COMM_NVLS"] = "1" if nvls else "0"

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world,
                            device_id=torch.device("cuda", rank))
    from sglang_security_b200.distributed import parallel_state as ps
    from sglang_security_b200.layers.linear import RowParallelLinear
    from sglang_security_b200.layers.quantization.w8a8_fp8 import W8A8Fp8Config

    ps.initialize_model_parallel(world)
    comm = ps.get_tp_custom_all_reduce()
    assert comm is not None and not comm.disabled
    g = torch.Generator().manual_seed(1234)  # same stream of numbers on every rank
    if nvls:
        _nvls_checks(rank, world, comm, ps, g, out_dir)
        dist.barrier()
        ps.destroy_model_parallel()
        dist.destroy_process_group()
        return
    for numel in [8, 4096, 64 * 4096, 256 * 8192]:
        xs = [(torch.randn(numel, generator=g) * 2).to(torch.bfloat16) for _ in range(world)]
        mine = xs[rank].cuda()
        nccl = mine.clone()  # taken first: above the crossover the group reduces `mine` in place
        out = ps.tensor_model_parallel_all_reduce(mine)
        if numel * 2 <= comm.crossover:  # above the measured crossover the group uses NCCL
            assert torch.equal(out.cpu(), OA.all_reduce_oracle(xs)), numel
        dist.all_reduce(nccl)
        # NCCL rounds to bf16 between its reduction steps: allow one bf16 ulp of the largest sums
        torch.testing.assert_close(out.float(), nccl.float(), rtol=2e-2, atol=2e-2 * world)
    # message above max_size falls back to NCCL
    big = torch.ones(comm.max_size // 2 + 64, dtype=torch.bfloat16, device="cuda")
    assert comm.custom_all_reduce(big) is None
    assert torch.all(ps.tensor_model_parallel_all_reduce(big) == world)

    # RowParallelLinear with the per-channel FP8 method: fused GEMM + all-reduce == separate
    hidden, inter = 1024, 4096
    w_full = torch.randn((hidden, inter), generator=g) * 0.02
    layer = RowParallelLinear(inter, hidden, bias=True, params_dtype=torch.bfloat16,
                              quant_config=W8A8Fp8Config(is_checkpoint_fp8_serialized=False)).cuda()
    layer.weight_loader(layer.weight, w_full.to(torch.bfloat16).cuda())
    layer.bias.data.copy_(torch.randn(hidden, generator=g).to(torch.bfloat16))
    layer.quant_method.process_weights_after_loading(layer)
    from sglang_security_b200.layers.quantization.fp8_kernel import fp8_gemm_smallm_config

    fp8_gemm_smallm_config(max_m=128)  # un-fused reference GEMM through the same swap-AB kernel
    from sglang_security_b200.layers.quantization.fp8_kernel import (comm_fuse_config,
                                                                    fp8_blockwise_scaled_mm)

    for fuse_max in (-1, 128):  # tuned policy, then fusion forced for every supported M
        comm_fuse_config(fuse_max)
        for m in [1, 16, 64, 128, 300]:
            x_full = torch.randn((m, inter), generator=g).to(torch.bfloat16)
            x = x_full[:, rank * (inter // world):(rank + 1) * (inter // world)].contiguous().cuda()
            y, _ = layer(x)                                         # GEMM + all-reduce in one call
            part = layer.quant_method.apply(layer, x, bias=layer.bias if rank == 0 else None)
            parts = [torch.empty_like(part) for _ in range(world)]
            dist.all_gather(parts, part)
            assert torch.equal(y.cpu(), OA.row_parallel_linear_oracle(parts)), (fuse_max, m)
            # an activation quantised by its producer (fused RMSNorm / SiLU*mul -> quant) takes the
            # same path and gives the same bytes
            from sglang_security_b200.layers.quantization.fp8_kernel import sglang_per_token_quant_fp8
            from sglang_security_b200.layers.quantization.fp8_utils import QuantizedActivation

            xq, xs = sglang_per_token_quant_fp8(x)
            yq, _ = layer(QuantizedActivation(xq, xs, torch.bfloat16))
            assert torch.equal(yq, y), ("quantised input", fuse_max, m)
        # block-scaled kernel: fused (M <= 64) and big tile + stand-alone all-reduce (M = 150)
        k = 1024
        wq = ((torch.rand((hidden, k), generator=g) - 0.5) * 8).to(FP8).cuda()
        ws = (torch.rand((hidden // 128, k // 128), generator=g) * 1e-2 + 1e-3).cuda()
        for m in [4, 64, 150]:
            aq = ((torch.rand((m, k), generator=g) - 0.5) * 16).to(FP8).cuda()
            a_s = (torch.rand((m, k // 128), generator=g) * 1e-2 + 1e-3).cuda()
            y = fp8_blockwise_scaled_mm(aq, wq, a_s, ws, torch.bfloat16, comm=comm)
            part = fp8_blockwise_scaled_mm(aq, wq, a_s, ws, torch.bfloat16)
            parts = [torch.empty_like(part) for _ in range(world)]
            dist.all_gather(parts, part)
            assert torch.equal(y.cpu(), OA.row_parallel_linear_oracle(parts)), ("block", fuse_max, m)
    comm_fuse_config(-1)
    dist.barrier()
    ps.destroy_model_parallel()
    open(os.path.join(out_dir, f"ok{rank}"), "w").write("ok")
    dist.destroy_process_group()


def _nvls_checks(rank, world, comm, ps, g, out_dir):
    """In-switch reduction (multimem.ld_reduce): fp32 accumulation inside the NVSwitch, one rounding,
    but in the switch's own order -> compared with the fp32 sum at one 16-bit ulp, with NCCL, and
    across ranks; CUDA-graph replays and both parities of the double buffer are exercised."""
    import torch.distributed as dist

    if not comm.nvls:   # fabric / driver / container without multicast support: recorded, not failed
        open(os.path.join(out_dir, f"nvls_unavailable{rank}"), "w").write("x")
        open(os.path.join(out_dir, f"ok{rank}"), "w").write("ok")
        return
    for dtype, ulp in ((torch.bfloat16, 2 ** -7), (torch.float16, 2 ** -10), (torch.float32, 2 ** -21)):
        for numel in [8192, 64 * 4096, 64 * 4096 + 8, 256 * 8192]:
            if numel * torch.empty(0, dtype=dtype).element_size() > comm.max_size:
                continue
            for rep in range(3):   # epochs of both parities
                xs = [(torch.randn(numel, generator=g) * 2).to(dtype) for _ in range(world)]
                mine = xs[rank].cuda()
                out = comm.all_reduce(mine)
                exact = torch.stack([x.double() for x in xs]).sum(0)
                err = (out.cpu().double() - exact).abs()
                # one rounding of the fp32 sum, but the switch does not round to nearest: measured up
                # to one full ulp (2-rank sums off by 0.055 at |x| ~ 8 in bf16)
                tol = ulp * exact.abs() * 1.01 + ulp * 2 * world
                assert (err <= tol).all(), (str(dtype), numel, rep, err.max().item())
                outs = [torch.empty_like(out) for _ in range(world)]
                dist.all_gather(outs, out)
                same = all(torch.equal(outs[0], o) for o in outs)
                if rank == 0 and not same:
                    open(os.path.join(out_dir, "nvls_ranks_differ"), "a").write(f"{dtype} {numel}\n")
    # CUDA graph: capture three all-reduces, replay twice, compare with an eager run
    x = (torch.randn(64 * 4096, generator=g) * 2).to(torch.bfloat16).cuda()
    ref = comm.all_reduce(comm.all_reduce(comm.all_reduce(x)))
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.stream(s):
        comm.all_reduce(x)  # warm-up on the side stream
        with torch.cuda.graph(graph, stream=s):
            y = comm.all_reduce(comm.all_reduce(comm.all_reduce(x)))
    torch.cuda.current_stream().wait_stream(s)
    for _ in range(2):
        graph.replay()
        torch.cuda.synchronize()
        torch.testing.assert_close(y.float(), ref.float(), rtol=2e-2, atol=1e-1 * world ** 3)
    open(os.path.join(out_dir, f"ok{rank}"), "w").write("ok")


@pytest.mark.parametrize("world", [2, 4, 8])
def test_nvls_all_reduce_multi_gpu(world, tmp_path):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    import torch.multiprocessing as mp

    mp.spawn(_mp_worker, args=(world, _free_port(), str(tmp_path), True), nprocs=world, join=True)
    assert all((tmp_path / f"ok{r}").exists() for r in range(world))
    if (tmp_path / "nvls_unavailable0").exists():
        pytest.skip("multicast objects are not available on this box (no NVLS)")


@pytest.mark.parametrize("world", [2, 4, 8])
def test_ipc_all_reduce_and_row_parallel_multi_gpu(world, tmp_path):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    import torch.multiprocessing as mp

    mp.spawn(_mp_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    assert all((tmp_path / f"ok{r}").exists() for r in range(world))
