"""Tensor-parallel all-reduce over NVLink peer memory (one process per GPU).

Host-side mirror of the reference's `CustomAllreduce`
(python/sglang/srt/distributed/device_communicators/custom_all_reduce.py:146-510): same life
cycle (`create_shared_buffer` = cudaMalloc + cudaIpc handle all-gathered through the process group,
:303-326), same predicates (`should_custom_ar` :414-435: byte size % 16, weakly contiguous,
below `max_size`) and the same `custom_all_reduce(input) -> Optional[Tensor]` contract
(:475-498: out-of-place, None = "use NCCL").  The kernels are ours (csrc/comm.cu one-shot push
all-reduce; csrc/fp8_gemm_smallm.cuh row-parallel GEMM with the reduce fused in its epilogue).
The epoch counters live in device memory, so captured CUDA graphs replay without re-registration
(`capture()` / `register_graph_buffers` of the reference are not needed).
"""
from __future__ import annotations

import ctypes
import os
import socket
import tempfile
import uuid
from typing import List, Optional

import torch
import torch.distributed as dist
from torch.distributed import ProcessGroup

from ... import _lib

# decode-sized messages only: bs 256 x hidden 8192 bf16 = 4 MiB is the largest row-parallel
# output of the named configs' decode sweeps; above that NCCL (ring/NVLS) wins on bandwidth.
_MAX_CAR_SIZE = 4 * 1024 * 1024


def is_weak_contiguous(inp: torch.Tensor) -> bool:
    return inp.is_contiguous() or (
        inp.untyped_storage().nbytes() - inp.storage_offset() * inp.element_size()
        == inp.numel() * inp.element_size())


class B200AllReduce:
    _SUPPORTED_WORLD_SIZES = [2, 3, 4, 5, 6, 7, 8]
    _SUPPORTED_DTYPES = (torch.bfloat16, torch.float16, torch.float32)

    def __init__(self, group: Optional[ProcessGroup], device: torch.device,
                 max_size: int = _MAX_CAR_SIZE) -> None:
        self.disabled = True
        self.handle = None
        self.group = group
        self.device = torch.device(device)
        self.rank = dist.get_rank(group=group)
        self.world_size = dist.get_world_size(group=group)
        if self.world_size == 1 or self.world_size not in self._SUPPORTED_WORLD_SIZES:
            return
        if self.device.type != "cuda":
            return
        lib = _lib.load()
        self.max_size = int(max_size)
        self._buf_bytes = int(lib.b200_comm_buffer_bytes(self.world_size, self.max_size))
        self.buffer_ptrs = self.create_shared_buffer(self._buf_bytes, group=group)
        arr = (ctypes.c_void_p * self.world_size)(*self.buffer_ptrs)
        h = ctypes.c_void_p()
        _lib.check(lib.b200_comm_create(self.rank, self.world_size, arr, self._buf_bytes,
                                        ctypes.byref(h)), "b200_comm_create")
        self.handle = h
        self.max_size = int(lib.b200_comm_max_bytes(h))
        # measured crossover with NCCL (NVLS) on 8 x B200 (profiles/r01_bench_allreduce_tp8.json): the
        # one-shot kernel is 24 vs 36 us at 1 MiB, equal at 2 MiB, 70 vs 61 us at 4 MiB; at TP = 2 it
        # stays ahead up to the 4 MiB buffer limit
        self.crossover = self.max_size if self.world_size <= 4 else min(self.max_size, 2 << 20)
        self.disabled = False
        dist.barrier(group=group)  # every rank has zeroed + mapped all buffers before first use
        # in-switch reduction (NVLS) for the messages it wins on; silently absent where the fabric,
        # the driver or the container does not offer multicast objects
        self.nvls = False
        self._nvls_buf = None
        # Default: on for groups of 8 (measured on 8 x B200, profiles/r02_bench_allreduce_tp8.json:
        # 15.9 vs 24.3 us at 512 KiB, whole-model TP 8 decode 3.86 vs 4.67 ms per step), off below
        # (TP 2: the LL kernel is 2x faster).  B200_COMM_NVLS=1 / 0 forces it on / off.
        want = os.environ.get("B200_COMM_NVLS", "")
        if want == "1" or (want != "0" and self.world_size >= 8):
            self._setup_nvls(group)
            if self.nvls:
                self.crossover = self.max_size  # the switch path keeps its lead up to the buffer limit

    # -- NVLS ------------------------------------------------------------------------------------
    def _all_ok(self, ok: bool, group) -> bool:
        flags = [None] * self.world_size
        dist.all_gather_object(flags, bool(ok), group=group)
        return all(flags)

    def _setup_nvls(self, group) -> None:
        """Multicast object shared by the group (csrc/comm_nvls.cu): created on rank 0, its file
        descriptor handed to the peers over a unix socket (SCM_RIGHTS), every rank adds its device,
        binds its own physical memory and maps both views.  Every step is agreed on by all ranks,
        so either all of them end up with NVLS or none does."""
        lib = _lib.load()
        size = 0
        if lib.b200_nvls_supported():
            size = int(lib.b200_nvls_buffer_bytes(self.world_size, 2 * self.max_size))
        sizes = [None] * self.world_size
        dist.all_gather_object(sizes, size, group=group)
        if not all(sizes) or len(set(sizes)) != 1:
            return
        buf = ctypes.c_void_p()
        path = [None]
        srv = None
        ok = True
        fd = ctypes.c_int(-1)
        if self.rank == 0:
            ok = lib.b200_nvls_create(self.world_size, size, ctypes.byref(buf), ctypes.byref(fd)) == 0
            if ok:
                try:   # a failure here must still reach the broadcast below (path stays None)
                    name = os.path.join(tempfile.gettempdir(), f"b200_nvls_{os.getpid()}_{uuid.uuid4().hex}")
                    srv = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
                    srv.bind(name)
                    srv.listen(self.world_size)
                    path[0] = name
                except OSError:
                    ok = False
        src = dist.get_global_rank(group, 0) if group is not None else 0
        dist.broadcast_object_list(path, src=src, group=group)
        try:
            if path[0] is None:
                ok = False
            elif self.rank == 0:
                srv.settimeout(60)
                for _ in range(self.world_size - 1):
                    conn, _addr = srv.accept()
                    socket.send_fds(conn, [b"fd"], [fd.value])
                    conn.close()
            else:
                cli = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
                cli.settimeout(60)
                cli.connect(path[0])
                _msg, fds, _flags, _addr = socket.recv_fds(cli, 16, 1)
                cli.close()
                ok = len(fds) == 1 and lib.b200_nvls_import(fds[0], self.world_size, size,
                                                            ctypes.byref(buf)) == 0
                for f in fds:
                    os.close(f)
        except Exception:  # noqa: BLE001 - any failure of the fd hand-off just means "no NVLS"
            ok = False
        finally:
            if srv is not None:
                srv.close()
                if path[0] is not None:
                    try:
                        os.unlink(path[0])
                    except OSError:
                        pass
            if fd.value >= 0:          # rank 0's exported descriptor: the peers hold their own copies
                os.close(fd.value)
        with torch.cuda.device(self.device):
            ok = self._all_ok(ok, group) and lib.b200_nvls_add_device(buf) == 0
            ok = self._all_ok(ok, group)   # also the barrier between add_device and bind
            uc, mc = ctypes.c_void_p(), ctypes.c_void_p()
            ok = ok and lib.b200_nvls_bind(buf, ctypes.byref(uc), ctypes.byref(mc)) == 0
            ok = self._all_ok(ok, group)
            if ok:
                ok = lib.b200_comm_attach_nvls(self.handle, buf) == 0
            ok = self._all_ok(ok, group)
        if ok:
            self.nvls, self._nvls_buf = True, buf
        elif buf.value:
            lib.b200_comm_attach_nvls(self.handle, None)
            lib.b200_nvls_free(buf)

    # -- buffers -------------------------------------------------------------------------------
    def create_shared_buffer(self, size_in_bytes: int,
                             group: Optional[ProcessGroup] = None) -> List[int]:
        """cudaMalloc a zero-filled buffer, exchange its IPC handle, open the peers' buffers;
        returns the device pointers of every rank's buffer in rank order."""
        lib = _lib.load()
        hb = lib.b200_comm_handle_bytes()
        handle = ctypes.create_string_buffer(hb)
        ptr = ctypes.c_void_p()
        with torch.cuda.device(self.device):
            _lib.check(lib.b200_comm_alloc(size_in_bytes, ctypes.byref(ptr), handle),
                       "b200_comm_alloc")
            handles = [None] * self.world_size
            dist.all_gather_object(handles, handle.raw, group=group)
            pointers: List[int] = []
            for i, h in enumerate(handles):
                if i == self.rank:
                    pointers.append(ptr.value)
                else:
                    p = ctypes.c_void_p()
                    _lib.check(lib.b200_comm_open(ctypes.create_string_buffer(h, hb),
                                                  ctypes.byref(p)), "b200_comm_open")
                    pointers.append(p.value)
        return pointers

    def free_shared_buffer(self) -> None:
        lib = _lib.load()
        for i, p in enumerate(self.buffer_ptrs):
            if i == self.rank:
                lib.b200_comm_free(ctypes.c_void_p(p))
            else:
                lib.b200_comm_close(ctypes.c_void_p(p))
        self.buffer_ptrs = []

    # -- predicates ----------------------------------------------------------------------------
    def should_custom_ar(self, inp: torch.Tensor) -> bool:
        if self.disabled or not inp.is_cuda or inp.dtype not in self._SUPPORTED_DTYPES:
            return False
        inp_size = inp.numel() * inp.element_size()
        if inp_size % 16 != 0 or inp.data_ptr() % 16 != 0:
            return False
        if not is_weak_contiguous(inp):
            return False
        return inp_size <= self.crossover

    # -- collectives ---------------------------------------------------------------------------
    def all_reduce(self, inp: torch.Tensor, *, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        if out is None:
            out = torch.empty_like(inp)
        _lib.check(_lib.load().b200_comm_all_reduce(
            self.handle, inp.data_ptr(), out.data_ptr(), inp.numel(), _lib.dtype_code(inp.dtype),
            _lib.current_stream()), "b200_comm_all_reduce")
        return out

    def custom_all_reduce(self, input: torch.Tensor) -> Optional[torch.Tensor]:
        """Out-of-place all-reduce, or None when the message must go through NCCL."""
        if self.disabled or not self.should_custom_ar(input):
            return None
        return self.all_reduce(input)

    def close(self) -> None:
        if not self.disabled and self.handle is not None:
            torch.cuda.synchronize(self.device)
            if dist.is_initialized():
                dist.barrier(group=self.group)
            _lib.load().b200_comm_destroy(self.handle)
            self.handle = None
            if self._nvls_buf is not None:
                _lib.load().b200_nvls_free(self._nvls_buf)
                self._nvls_buf, self.nvls = None, False
            self.free_shared_buffer()
            self.disabled = True


class LocalCommGroup:
    """`world` communicators that live in ONE process on ONE device (every "rank" is driven from
    its own CUDA stream).  Test/bench harness for the protocol on a single-GPU box: the kernels
    are exactly the ones the multi-process path launches, only the buffers are not IPC-mapped."""

    def __init__(self, world: int, max_size: int = _MAX_CAR_SIZE, device="cuda"):
        lib = _lib.load()
        self.world = world
        self.device = torch.device(device)
        nbytes = int(lib.b200_comm_buffer_bytes(world, max_size))
        self._bufs = [torch.zeros(nbytes, dtype=torch.uint8, device=self.device)
                      for _ in range(world)]
        arr = (ctypes.c_void_p * world)(*[b.data_ptr() for b in self._bufs])
        self.comms = []
        for r in range(world):
            h = ctypes.c_void_p()
            _lib.check(lib.b200_comm_create(r, world, arr, nbytes, ctypes.byref(h)),
                       "b200_comm_create")
            self.comms.append(_LocalComm(h, r, world))
        self.streams = [torch.cuda.Stream(device=self.device) for _ in range(world)]
        torch.cuda.synchronize(self.device)

    def close(self):
        torch.cuda.synchronize(self.device)
        for c in self.comms:
            _lib.load().b200_comm_destroy(c.handle)
        self.comms = []


class _LocalComm:
    def __init__(self, handle, rank, world):
        self.handle, self.rank, self.world_size, self.disabled = handle, rank, world, False

    def all_reduce(self, inp, *, out=None):
        if out is None:
            out = torch.empty_like(inp)
        _lib.check(_lib.load().b200_comm_all_reduce(
            self.handle, inp.data_ptr(), out.data_ptr(), inp.numel(), _lib.dtype_code(inp.dtype),
            _lib.current_stream()), "b200_comm_all_reduce")
        return out
