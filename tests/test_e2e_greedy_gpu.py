"""BASELINE config 1 end to end: Qwen2.5-0.5B geometry, random-init bf16 weights, a 512-token shared
radix prefix, 8 requests x 16 own tokens, 32 greedy decode steps — the B200 path (attention backend
+ linear layers through the C-ABI) against the CPU oracle stack on the same weights."""
import pytest
import torch

import model_stack as MS

pytestmark = pytest.mark.gpu


def _session(cfg, decode_steps=32, bs=8, prefix=512, own=16):
    torch.manual_seed(0)
    weights = MS.init_weights(cfg, seed=0, std=0.02)
    g = torch.Generator().manual_seed(1)
    prefix_ids = torch.randint(0, cfg.vocab, (prefix,), generator=g)
    suffix = [torch.randint(0, cfg.vocab, (own,), generator=g) for _ in range(bs)]
    kw = dict(pool_size=prefix + bs * (own + decode_steps) + 8, max_reqs=bs + 1,
              max_ctx=prefix + own + decode_steps + 4)
    ref = MS.Stack(cfg, weights, "oracle", **kw)
    tok_ref, logits_ref = MS.run_shared_prefix_session(ref, prefix_ids, suffix, decode_steps)
    gpu = MS.Stack(cfg, weights, "b200", **kw)
    # teacher-forced with the oracle's tokens so every step is compared on identical inputs
    tok_gpu, logits_gpu = MS.run_shared_prefix_session(gpu, prefix_ids, suffix, decode_steps,
                                                       forced_tokens=tok_ref)
    return ref, gpu, tok_ref, tok_gpu, logits_ref, logits_gpu


def _check_streams(tok_ref, tok_gpu, logits_ref, logits_gpu, logit_tol):
    """Teacher-forced comparison: at every step both stacks see identical inputs.  Logits must
    agree within logit_tol; a differing greedy token is accepted only when the oracle's own logit
    gap between the two candidates is within twice the logit error measured at that step (a tie
    the measured rounding noise can flip) — otherwise the streams must be identical."""
    mism, worst, total = 0, 0.0, 0
    for step, (lr, lg) in enumerate(zip(logits_ref, logits_gpu)):
        step_worst = (lr - lg).abs().max().item()
        worst = max(worst, step_worst)
        bad = (tok_ref[step] != tok_gpu[step]).nonzero().flatten()
        total += tok_ref[step].numel()
        for i in bad.tolist():
            gap = (lr[i, tok_ref[step][i]] - lr[i, tok_gpu[step][i]]).item()
            assert 0 <= gap <= 2 * step_worst, (step, i, gap, step_worst)
        mism += len(bad)
    assert worst <= logit_tol, worst
    print(f"[e2e] {total - mism}/{total} greedy tokens identical, {mism} tie flips, "
          f"worst logit diff {worst:.4f}")
    return mism, worst


def test_greedy_streams_match_bf16_qwen05b_geometry():
    cfg = MS.StackConfig()  # Qwen2.5-0.5B: hidden 896, 14/2 heads, D 64, 24 layers
    ref, gpu, tok_ref, tok_gpu, lr, lg = _session(cfg)
    # 24 bf16 layers, CPU (oneDNN) vs GPU (cuBLAS) accumulation order: logits (|x| up to ~4, bf16
    # ulp 1.6e-2..3.1e-2) measured to agree within 0.073 on B200; tolerance 0.1 = ~4 ulp
    mism, worst = _check_streams(tok_ref, tok_gpu, lr, lg, logit_tol=1e-1)
    assert mism <= 0.1 * tok_ref.numel(), (mism, worst)
    # KV pools: shared-prefix rows and every written row agree to bf16 tolerance in all layers,
    # req_to_token is identical
    assert torch.equal(gpu.rtp.req_to_token.cpu(), ref.rtp.req_to_token)
    used = ref.next_slot
    for l in (0, cfg.layers // 2, cfg.layers - 1):
        dk = (gpu.kvp.k_buffer[l][:used].float().cpu() - ref.kvp.k_buffer[l][:used].float()).abs().max()
        assert dk.item() <= 0.25, (l, dk)


def test_greedy_streams_match_fp8_llama_geometry_small():
    # Llama-style GQA group 4, D 128, FP8 W8A8 per-channel linear layers (cfg 3 path), 4 layers
    cfg = MS.StackConfig(hidden=1024, heads=8, kv_heads=2, head_dim=128, inter=2816, layers=4,
                         vocab=8192, qkv_bias=False, quant="fp8")
    ref, gpu, tok_ref, tok_gpu, lr, lg = _session(cfg, decode_steps=8, bs=4, prefix=200, own=9)
    # FP8 re-quantisation amplifies bf16-ulp differences between cuBLAS/oneDNN and between the
    # tcgen05 and CPU accumulation orders: perturbing the ORACLE's own embeddings by 0.2 % (half a
    # bf16 ulp) moves its logits (std 0.64) by 0.23, the B200 path differs from the oracle by 0.24.
    # Random-init logits are nearly tied (top-2 gap of 8192 N(0, 0.64) draws ~0.15 vs 0.24 of
    # measured noise), so tie flips are expected; _check_streams already asserts that EVERY flip is
    # a tie inside the measured noise.  Measured on B200: 8 flips of 36 tokens; cap at one third.
    mism, worst = _check_streams(tok_ref, tok_gpu, lr, lg, logit_tol=3.5e-1)
    assert mism <= tok_ref.numel() // 3, (mism, worst)


@pytest.mark.parametrize("quant", [None, "fp8"])
def test_fused_layer_equals_unfused_layer_bit_exact(quant):
    """B200DecoderLayer(fuse=True) — RMSNorm->quant, RoPE->KV write, SiLU*mul->quant fused — must
    give exactly the logits and KV pool of the op-by-op layer in the reference's call order."""
    cfg = MS.StackConfig(hidden=1024, heads=8, kv_heads=2, head_dim=128, inter=2816, layers=3,
                         vocab=4096, qkv_bias=quant is None, quant=quant)
    weights = MS.init_weights(cfg, seed=3, std=0.02)
    g = torch.Generator().manual_seed(5)
    prefix_ids = torch.randint(0, cfg.vocab, (70,), generator=g)
    suffix = [torch.randint(0, cfg.vocab, (n,), generator=g) for n in (5, 9, 1)]
    kw = dict(pool_size=200, max_reqs=5, max_ctx=100)
    outs = []
    for fuse in (True, False):
        st = MS.Stack(cfg, weights, "b200", fuse=fuse, **kw)
        tok, logits = MS.run_shared_prefix_session(st, prefix_ids, suffix, 4)
        outs.append((tok, logits, st))
    assert torch.equal(outs[0][0], outs[1][0])
    for a, b in zip(outs[0][1], outs[1][1]):
        assert torch.equal(a, b)
    for l in range(cfg.layers):
        assert torch.equal(outs[0][2].kvp.k_buffer[l], outs[1][2].kvp.k_buffer[l])
        assert torch.equal(outs[0][2].kvp.v_buffer[l], outs[1][2].kvp.v_buffer[l])


def _free_running(cfg, embed_std, decode_steps, bs, prefix, own, noise_budget):
    """Both stacks decode FREE-RUNNING (each feeds back its own argmax).  The init gives every greedy
    decision a margin far above the numeric noise (model_stack.init_weights, decisive_embed_std), so
    the token streams must be IDENTICAL; the margin / noise ratio is measured and asserted."""
    weights = MS.init_weights(cfg, seed=0, std=0.02, decisive_embed_std=embed_std)
    g = torch.Generator().manual_seed(1)
    prefix_ids = torch.randint(0, cfg.vocab, (prefix,), generator=g)
    suffix = [torch.randint(0, cfg.vocab, (own,), generator=g) for _ in range(bs)]
    kw = dict(pool_size=prefix + bs * (own + decode_steps) + 8, max_reqs=bs + 1,
              max_ctx=prefix + own + decode_steps + 4)
    ref = MS.Stack(cfg, weights, "oracle", **kw)
    tok_ref, logits_ref = MS.run_shared_prefix_session(ref, prefix_ids, suffix, decode_steps)
    gpu = MS.Stack(cfg, weights, "b200", **kw)
    tok_gpu, logits_gpu = MS.run_shared_prefix_session(gpu, prefix_ids, suffix, decode_steps)
    min_margin, worst = float("inf"), 0.0
    for lr, lg in zip(logits_ref, logits_gpu):
        top = lr.topk(2, dim=-1).values
        min_margin = min(min_margin, (top[:, 0] - top[:, 1]).min().item())
        worst = max(worst, (lr - lg).abs().max().item())
    sigma = logits_ref[0].std().item()
    print(f"[e2e free-running] {tok_ref.numel()} tokens, min top-1 margin {min_margin:.1f}, worst "
          f"logit diff {worst:.2f} (logit std {sigma:.1f}) -> margin/noise {min_margin / max(worst, 1e-9):.1f}")
    assert torch.equal(tok_ref, tok_gpu), (tok_ref != tok_gpu).sum().item()
    assert worst <= noise_budget * sigma, (worst, sigma)
    assert min_margin >= 4 * worst, (min_margin, worst)
    # the stream is not degenerate: requests produce different, changing tokens
    assert tok_ref.unique().numel() >= tok_ref.numel() // 2
    assert torch.equal(gpu.rtp.req_to_token.cpu(), ref.rtp.req_to_token)
    return ref, gpu


def test_greedy_free_running_identical_bf16_cfg1():
    """BASELINE cfg 1 at full geometry (Qwen2.5-0.5B: 24 layers, hidden 896, 14/2 heads, D 64, qkv
    bias, tied head), 512-token shared radix prefix, 8 requests x 16 own tokens, 32 greedy decode
    steps: 264 greedy tokens, zero mismatches."""
    cfg = MS.StackConfig()
    ref, gpu = _free_running(cfg, 1.0, decode_steps=32, bs=8, prefix=512, own=16, noise_budget=0.15)
    used = ref.next_slot
    for l in (0, cfg.layers - 1):
        dk = (gpu.kvp.k_buffer[l][:used].float().cpu() - ref.kvp.k_buffer[l][:used].float()).abs().max()
        assert dk.item() <= 0.5, (l, dk)


def test_greedy_free_running_identical_fp8_small():
    cfg = MS.StackConfig(hidden=1024, heads=8, kv_heads=2, head_dim=128, inter=2816, layers=4,
                         vocab=8192, qkv_bias=False, quant="fp8")
    _free_running(cfg, 1.0, decode_steps=16, bs=4, prefix=200, own=9, noise_budget=0.5)


def test_greedy_free_running_identical_fp8_llama8b_width():
    """FP8 W8A8 linear layers at the full Llama-3.1-8B layer width (hidden 4096, 32/8 heads, D 128,
    intermediate 14336), 2 layers."""
    cfg = MS.StackConfig(hidden=4096, heads=32, kv_heads=8, head_dim=128, inter=14336, layers=2,
                         vocab=16384, rope_theta=5e5, eps=1e-5, qkv_bias=False, quant="fp8")
    _free_running(cfg, 1.0, decode_steps=8, bs=4, prefix=128, own=8, noise_budget=0.5)
