"""b200_attn_decode_noalloc (split-KV scratch built into the library, no device allocation) through
the C-ABI against the existing entry point b200_attn_decode (caller-provided workspace): same
kernels, so the outputs must agree within the decode parity tolerance (max-abs 3e-2, rel-l2 1e-2),
on two calls in a row with different shapes (the second call finds the first call's partials in the
shared scratch) and on the single-split path; requests beyond the 8 MiB scratch must be rejected.
Collected last on purpose (file name): a failure here must not hide other tests under `pytest -x`."""
import pytest
import torch

from oracle import scenarios

pytestmark = pytest.mark.gpu
SCRATCH_BYTES = 8 << 20   # B200_ATTN_DECODE_NOALLOC_SCRATCH_BYTES (include/b200_hotpath.h)


def _device_case(seed, hq, hkv, d, seq_lens, dtype):
    case = scenarios.make_attention_case(seed=seed, hq=hq, hkv=hkv, d=d, seq_lens=seq_lens,
                                         shared_prefix=16, dtype=dtype)
    kc, vc = case.k_cache.clone(), case.v_cache.clone()
    kc[case.out_cache_loc] = case.k_new
    vc[case.out_cache_loc] = case.v_new
    dev = "cuda"
    return dict(hq=hq, hkv=hkv, d=d, bs=case.bs, dtype=dtype,
                q=case.q.to(dev).contiguous(), k=kc.to(dev).reshape(kc.shape[0], -1).contiguous(),
                v=vc.to(dev).reshape(vc.shape[0], -1).contiguous(),
                rtt=case.req_to_token.to(dev).contiguous(),
                rpi=case.req_pool_indices.to(dev).to(torch.int64),
                lens=case.seq_lens.to(dev).to(torch.int64))


def _args(c, o):
    from sglang_security_b200 import _lib

    return (c["q"].data_ptr(), c["q"].stride(0), c["k"].data_ptr(), c["v"].data_ptr(), c["k"].stride(0),
            o.data_ptr(), o.stride(0), c["rtt"].data_ptr(), c["rtt"].stride(0), c["rpi"].data_ptr(),
            c["lens"].data_ptr(), c["bs"], c["hq"], c["hkv"], c["d"], c["d"], float(c["d"]) ** -0.5, 0.0,
            _lib.dtype_code(c["dtype"]))


def _run_both(c, splits):
    """-> (o from b200_attn_decode with a torch workspace, o from b200_attn_decode_noalloc)"""
    from sglang_security_b200 import _lib

    lib = _lib.load()
    stream = _lib.current_stream()
    need = lib.b200_attn_decode_workspace_bytes(c["bs"], c["hq"], c["d"], splits)
    assert need <= SCRATCH_BYTES
    ws = torch.empty(max(need, 16), dtype=torch.uint8, device="cuda")
    o_ref = torch.zeros((c["bs"], c["hq"] * c["d"]), dtype=c["dtype"], device="cuda")
    o_new = torch.full_like(o_ref, float("nan"))
    _lib.check(lib.b200_attn_decode(*_args(c, o_ref), ws.data_ptr(), ws.numel(), splits, stream),
               "b200_attn_decode")
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    _lib.check(lib.b200_attn_decode_noalloc(*_args(c, o_new), splits, stream), "b200_attn_decode_noalloc")
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before
    return o_ref.float().cpu(), o_new.float().cpu()


def _close(o_ref, o_new):
    assert torch.isfinite(o_new).all()
    diff = (o_new - o_ref).abs()
    rel_l2 = float(diff.norm() / o_ref.norm().clamp_min(1e-12))
    assert float(diff.max()) <= 3e-2 and rel_l2 <= 1e-2, (float(diff.max()), rel_l2)


def test_noalloc_decode_matches_existing_entry_point_twice_in_a_row():
    a = _device_case(301, 32, 8, 128, [1, 63, 300, 1025, 700, 129], torch.bfloat16)
    b = _device_case(302, 14, 2, 64, [513, 3, 96], torch.float16)
    _close(*_run_both(a, 4))        # first call
    _close(*_run_both(b, 3))        # second call: other shapes over the first call's stale partials
    _close(*_run_both(a, 4))        # and the first shape again
    _close(*_run_both(a, 1))        # single split: no scratch use at all


def test_noalloc_decode_rejects_requests_beyond_the_scratch():
    from sglang_security_b200 import _lib

    lib = _lib.load()
    c = _device_case(303, 32, 8, 128, [40] * 64, torch.bfloat16)
    o = torch.zeros((c["bs"], c["hq"] * c["d"]), dtype=c["dtype"], device="cuda")
    assert lib.b200_attn_decode_workspace_bytes(64, 32, 128, 8) > SCRATCH_BYTES
    rc = lib.b200_attn_decode_noalloc(*_args(c, o), 8, _lib.current_stream())
    assert rc != 0
    msg = lib.b200_last_error().decode()
    assert "scratch" in msg and "8388608" in msg, msg
    torch.cuda.synchronize()
    assert not o.any()              # nothing was launched
    # the largest split count that fits is accepted
    _lib.check(lib.b200_attn_decode_noalloc(*_args(c, o), 7, _lib.current_stream()), "noalloc")
    torch.cuda.synchronize()
    assert o.any()
