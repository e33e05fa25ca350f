"""CPU: the oracle restatements of the quantised-linear ops against fixtures produced by the
reference's OWN in-test torch functions (oracle/gen_golden_quant.py extracts their source unmodified
from /root/reference and runs it).  This is what pins oracle/fp8.py and oracle/int4.py (AWQ)."""
import os

import torch

from oracle import fp8 as OF
from oracle import int4 as OI

GOLD = os.path.join(os.path.dirname(__file__), "golden")


def _load(name):
    return torch.load(os.path.join(GOLD, name + ".pt"), weights_only=False)


def zero_group_mask(x, group):
    return x.float().reshape(x.shape[0], -1, group).abs().amax(-1) == 0


def test_awq_dequant_oracle_equals_reference_function():
    for c in _load("awq_dequant"):
        out = OI.awq_dequantize_oracle(c["qweight"], c["scales"], c["qzeros"], c["group"])
        assert out.dtype == c["out"].dtype
        assert torch.equal(out.view(torch.int16), c["out"].view(torch.int16))
        # the integer unpack itself (bit-exact target of north_star)
        w = OI.unpack_awq_int4(c["qweight"])
        assert w.min() >= 0 and w.max() <= 15


def test_fp8_quantisers_oracle_equals_reference_functions():
    for c in _load("fp8_quant"):
        q, s = OF.per_token_quant_fp8_oracle(c["x"])
        assert torch.equal(s.reshape(-1), c["token_scale"])
        keep = torch.ones(c["x"].shape[0], dtype=torch.bool)
        keep[c["zero_rows"]] = False      # scale 0: the reference twin divides by it
        assert torch.equal(q[keep].view(torch.uint8), c["q_token"][keep].view(torch.uint8))
        assert (q[~keep].float() == 0).all()
        qg, sg = OF.per_token_group_quant_fp8_oracle(c["x_group"], 128)
        live = zero_group_mask(c["x_group"], 128).logical_not()
        assert torch.equal(sg[live], c["s_group"][live])
        live_q = live.repeat_interleave(128, dim=1)
        assert torch.equal(qg.view(torch.uint8)[live_q], c["q_group"].view(torch.uint8)[live_q])
        # all-zero groups: the reference's KERNELS clamp the fp32 absmax with eps
        # (fp8_kernel.py:133 `tl.maximum(tl.max(tl.abs(y)), eps)`, per_token_group_quant_8bit.cu:40
        # `local_absmax = eps`), its in-test torch twin clamps in the INPUT dtype (bf16 rounds
        # 1e-10, fp16 flushes it to 0 and then divides 0/0); the oracle follows the kernels
        assert (sg[~live] == torch.tensor(1e-10, dtype=torch.float32) / 448.0).all()
        assert (qg.float()[~live_q] == 0).all()
        qs = OF.static_quant_fp8_oracle(c["x"], c["static_scale"])
        assert torch.equal(qs.view(torch.uint8), c["q_static"].view(torch.uint8))


def test_fp8_scaled_mm_oracle_equals_reference_function():
    for c in _load("fp8_gemm"):
        out = OF.scaled_mm_oracle(c["a"], c["b"].t(), c["scale_a"], c["scale_b"], c["out"].dtype,
                                  c["bias"])
        assert torch.equal(out, c["out"])


def test_fp8_block_mm_oracle_equals_reference_function():
    for c in _load("fp8_block"):
        out = OF.block_scaled_mm_oracle(c["a"], c["b"], c["a_s"], c["b_s"], (128, 128),
                                        c["out"].dtype)
        # same tile order and fp32 accumulation as native_w8a8_block_fp8_matmul
        assert torch.equal(out, c["out"])


def test_gemv_stream_layout_oracle_against_scalar_loops():
    """The vectorised restatement of the decode GEMV's stream layout (our own format) against plain
    loops written from the format description in csrc/w4a16_gemv2.cu."""
    import numpy as np
    import torch

    from oracle import int4 as O

    k, n = 128, 32
    rs = np.random.RandomState(5)
    qweight = torch.from_numpy(rs.randint(-2**31, 2**31, size=(k, n // 8), dtype=np.int64).astype(np.int32))
    ints = O.unpack_awq_int4(qweight).numpy()
    got = O.gemv_stream_layout_oracle(qweight).numpy().astype(np.int64) & 0xFFFFFFFF
    order = [0, 2, 4, 6, 1, 3, 5, 7]
    idx = 0
    for nt in range(n // 16):
        for kb in range(k // 64):
            for lane in range(32):
                g, t = lane >> 2, lane & 3
                for j in range(4):
                    col = nt * 16 + g + 8 * (j & 1)
                    k0 = kb * 64 + (j >> 1) * 32 + 8 * t
                    w = 0
                    for i in range(8):
                        w |= int(ints[k0 + order[i], col]) << (4 * i)
                    assert got[idx] == w, (nt, kb, lane, j)
                    idx += 1
    assert idx == got.size


def test_gptq_qweight_packing_matches_vllm_gptq_pack():
    """Packing order of the GPTQ qweight tensor against vllm's own gptq_pack (the dependency the
    reference delegates GPTQ to; fixture from oracle/gen_golden_gptq_pack.py).  The dequant
    arithmetic has no CPU implementation in vllm: it is checked on the GPU (test_gptq_vs_vllm_gpu.py)."""
    import os

    import torch

    from oracle import int4 as O

    gold = torch.load(os.path.join(os.path.dirname(__file__), "golden", "gptq_pack_vllm.pt"))
    assert gold["vllm_version"]
    for c in gold["cases"]:
        assert torch.equal(O.unpack_gptq_qweight(c["qweight"]).to(torch.int32), c["q"])


def test_fp8_pool_write_oracle_equals_reference_method():
    """oracle/pools.py set_kv_buffer on e4m3 and e5m2 pools against the bytes the reference's own
    MHATokenToKVPool.set_kv_buffer wrote (memory_pool.py:378-412, extracted and run by
    oracle/gen_golden_kv_write.py): div_ by the scales, cast, byte view, index_put."""
    from types import SimpleNamespace

    from oracle.pools import OracleMHATokenToKVPool

    gold = _load("kv_write_fp8")
    assert len(gold["cases"]) == 12
    for c in gold["cases"]:
        pool_dtype = getattr(torch, c["pool_dtype"].split(".")[-1])
        pool = OracleMHATokenToKVPool(100, 1, pool_dtype, c["k"].shape[1], c["k"].shape[2], 1)
        ks, vs = c["scales"] if c["scales"] else (None, None)
        pool.set_kv_buffer(SimpleNamespace(layer_id=0), c["loc"], c["k"].clone(), c["v"].clone(), ks, vs)
        assert torch.equal(pool.k_buffer[0][c["loc"]], c["k_rows"]), (c["pool_dtype"], c["dtype"], c["scales"])
        assert torch.equal(pool.v_buffer[0][c["loc"]], c["v_rows"])
        assert pool.get_key_buffer(0).dtype == pool_dtype
