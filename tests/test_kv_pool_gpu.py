"""GPU parity (bit-exact) of the KV-pool write and the req_to_token gather."""
import numpy as np
import pytest
import torch

from oracle.pools import kv_indices_oracle

pytestmark = pytest.mark.gpu


def _kv_indices_cuda(req_to_token, req_pool_indices, lens, kv_start=None):
    from sglang_security_b200 import _lib

    lib = _lib.load()
    bs = req_pool_indices.numel()
    indptr = torch.zeros(bs + 1, dtype=torch.int32, device="cuda")
    indptr[1:] = torch.cumsum(lens, 0)
    out = torch.full((int(lens.sum().item()) + 8,), -7, dtype=torch.int32, device="cuda")
    rc = lib.b200_kv_indices(req_to_token.data_ptr(), req_to_token.stride(0),
                             req_pool_indices.data_ptr(), lens.data_ptr(), indptr.data_ptr(),
                             _lib.ptr(kv_start), out.data_ptr(), bs, _lib.current_stream())
    _lib.check(rc, "b200_kv_indices")
    torch.cuda.synchronize()
    return indptr, out


@pytest.mark.parametrize("batch", [1, 37, 1786])
def test_kv_indices_reference_shapes(batch):
    # test/srt/test_create_kvindices.py:18-71: 4096 x 4096 table, batch in {1, 37, 1786}
    rs = np.random.RandomState(batch)
    max_batch = max_ctx = 4096
    table = torch.arange(max_batch * max_ctx, dtype=torch.int32).reshape(max_batch, max_ctx)
    req = torch.from_numpy(rs.choice(max_batch, size=batch, replace=False).astype(np.int64))
    lens = torch.from_numpy(rs.choice(max_ctx, size=batch, replace=False).astype(np.int64))
    indptr_ref, flat_ref = kv_indices_oracle(table, req, lens)
    indptr, out = _kv_indices_cuda(table.cuda(), req.cuda(), lens.cuda())
    n = flat_ref.numel()
    assert torch.equal(indptr.cpu(), indptr_ref)
    assert torch.equal(out[:n].cpu(), flat_ref)
    assert (out[n:] == -7).all()  # nothing written past the end


def test_kv_indices_with_start_and_empty_rows():
    rs = np.random.RandomState(3)
    table = torch.from_numpy(rs.randint(0, 1 << 30, size=(64, 700)).astype(np.int32))
    req = torch.tensor([5, 9, 63, 0], dtype=torch.int64)
    lens = torch.tensor([0, 513, 1, 300], dtype=torch.int64)
    start = torch.tensor([10, 100, 699, 0], dtype=torch.int32)
    indptr_ref, flat_ref = kv_indices_oracle(table, req, lens, start)
    _, out = _kv_indices_cuda(table.cuda(), req.cuda(), lens.cuda(), start.cuda())
    assert torch.equal(out[:flat_ref.numel()].cpu(), flat_ref)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.uint8])
@pytest.mark.parametrize("hkv,d,tokens", [(8, 128, 64), (2, 64, 1), (1, 128, 777), (3, 80, 5)])
def test_kv_write_bit_exact(dtype, hkv, d, tokens):
    from sglang_security_b200.mem_cache.memory_pool import kv_write

    g = torch.Generator().manual_seed(tokens * 31 + d)
    size = 2048
    raw = torch.randint(0, 256, (size + 1, hkv, d * 2), generator=g, dtype=torch.uint8)
    es = torch.empty((), dtype=dtype).element_size()
    k_cache = raw[..., :d * es].contiguous().view(dtype).clone()
    v_cache = raw.flip(0)[..., :d * es].contiguous().view(dtype).clone()
    # new rows as strided slices of a fused qkv buffer
    fused = torch.randint(0, 256, (tokens, (4 + 2) * hkv * d * es), generator=g, dtype=torch.uint8).view(dtype)
    k = fused[:, 4 * hkv * d:5 * hkv * d].view(tokens, hkv, d)
    v = fused[:, 5 * hkv * d:].view(tokens, hkv, d)
    loc = (torch.randperm(size, generator=g)[:tokens] + 1).to(torch.int64)
    k_ref, v_ref = k_cache.clone(), v_cache.clone()
    k_ref[loc] = k
    v_ref[loc] = v
    kc, vc = k_cache.cuda(), v_cache.cuda()
    kv_write(k.cuda(), v.cuda(), kc, vc, loc.cuda())
    torch.cuda.synchronize()
    assert torch.equal(kc.cpu().view(torch.uint8), k_ref.view(torch.uint8))
    assert torch.equal(vc.cpu().view(torch.uint8), v_ref.view(torch.uint8))


def test_kv_write_padded_tokens_go_to_dummy_row():
    from sglang_security_b200.mem_cache.memory_pool import kv_write

    kc = torch.zeros(10, 2, 64, dtype=torch.bfloat16, device="cuda")
    vc = torch.zeros_like(kc)
    k = torch.ones(3, 2, 64, dtype=torch.bfloat16, device="cuda")
    kv_write(k, k * 2, kc, vc, torch.tensor([4, 0, 0], device="cuda"))
    torch.cuda.synchronize()
    assert (kc[4] == 1).all() and (vc[4] == 2).all()
    assert (kc[1:4] == 0).all() and (kc[5:] == 0).all()
