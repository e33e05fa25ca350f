"""B200AttnBackend — the single new AttentionBackend of the hot path.

Drop-in for the reference backends (TorchNativeAttnBackend is the numerical oracle,
python/sglang/srt/layers/attention/torch_native_backend.py:17-270; TritonAttnBackend the GPU
implementation of the same interface, triton_backend.py:151-835).  It indexes
`req_to_token_pool.req_to_token` and the `MHATokenToKVPool` buffers in place (layouts unchanged)
and calls the sm_100a kernels through the C-ABI (include/b200_hotpath.h):

  forward_decode : b200_kv_write -> b200_attn_decode_fused (split-KV gather kernel whose last CTA per
                   request merges the splits and, on request, quantises the row for o_proj);
                   B200_ATTN_FUSED_TAIL=0 restores split kernel + merge kernel
  forward_extend : b200_kv_write -> b200_attn_extend (tcgen05 prefix rectangle + causal triangle)

All per-step metadata is computed without a device->host sync; buffers are persistent so the
decode step can be captured in a CUDA graph.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, Optional

import os

import torch

from ... import _lib
from ...mem_cache.memory_pool import kv_write
from .base_attn_backend import AttentionBackend


@dataclass
class ForwardMetadata:
    num_splits: int = 1
    workspace: Optional[torch.Tensor] = None
    # extend only
    qo_indptr: Optional[torch.Tensor] = None
    max_extend_len: int = 0
    total_extend: int = 0
    # dtype-converted copies of the ForwardBatch index tensors, kept alive for the whole forward
    # (one conversion per step, never a temporary whose block the allocator could hand out again
    # before the launch has read it); keyed by (data_ptr, dtype, numel) of the source
    converted: Dict[tuple, torch.Tensor] = field(default_factory=dict)

    def as_dtype(self, t: torch.Tensor, dtype: torch.dtype) -> torch.Tensor:
        if t.dtype == dtype:
            return t
        key = (t.data_ptr(), t.dtype, dtype, t.numel())
        out = self.converted.get(key)
        if out is None:
            out = t.to(dtype)
            self.converted[key] = out
        return out


class B200AttnBackend(AttentionBackend):
    def __init__(self, model_runner, skip_prefill: bool = False):
        super().__init__()
        self.lib = _lib.load()  # raises if the CUDA library has not been built
        self.device = model_runner.device
        self.req_to_token = model_runner.req_to_token_pool.req_to_token
        tp = getattr(model_runner, "tp_size", 1)
        cfg = model_runner.model_config
        self.num_head = cfg.num_attention_heads // tp
        self.num_kv_head = cfg.get_num_kv_heads(tp)
        self.max_context_len = cfg.context_len
        self.skip_prefill = skip_prefill
        server_args = getattr(model_runner, "server_args", None)
        # 0 -> heuristic; otherwise a fixed split count (same knob as
        # --triton-attention-num-kv-splits, server_args.py:195)
        self.fixed_num_kv_splits = int(getattr(server_args, "b200_attention_num_kv_splits", 0) or 0)
        self.v_head_dim = model_runner.token_to_kv_pool.get_value_buffer(0).shape[-1]
        self.num_sms = torch.cuda.get_device_properties(self.device).multi_processor_count
        self.forward_metadata: Optional[ForwardMetadata] = None
        self._workspace: Optional[torch.Tensor] = None
        self._graph_workspace: Optional[torch.Tensor] = None
        self._graph_max_bs = 0
        # fused decode tail (merge + optional o_proj input quantisation by the last CTA of a request)
        # fused_tail: False = split + merge kernels and a separate quantiser (the reference's launch
        # structure).  True = ONE merge+quantise launch after the split kernel (a cluster of CTAs per
        # request).  The in-kernel variant (ticket counters inside the split kernel, batch <=
        # tail_in_kernel_max_bs) measured slower than both at every batch on B200
        # (profiles/r02_decode_tail_modes.json) and is therefore off by default.
        self.fused_tail = os.environ.get("B200_ATTN_FUSED_TAIL", "1") != "0"
        self.tail_in_kernel_max_bs = int(os.environ.get("B200_ATTN_TAIL_IN_KERNEL_MAX_BS", "0"))
        self._tickets = torch.zeros(256 * (1 + self.num_head), dtype=torch.int32, device=self.device)

    def _ticket_buffer(self, bs: int) -> torch.Tensor:
        """Zero-initialised, self-re-arming counters of the fused tail (one per request and one per
        (request, kv-head tile)); bs * (1 + num_head) words always cover the kernel's need."""
        need = bs * (1 + self.num_head)
        if self._tickets.numel() < need:
            if torch.cuda.is_current_stream_capturing():
                raise _lib.B200Error("B200AttnBackend: ticket buffer too small during graph capture "
                                     "(call init_cuda_graph_state with the largest batch first)")
            self._tickets = torch.zeros(max(need, 2 * self._tickets.numel()), dtype=torch.int32,
                                        device=self.device)
        return self._tickets

    # ------------------------------------------------------------------ metadata
    def _num_splits(self, bs: int, head_dim: int, avg_len: int) -> int:
        if self.fixed_num_kv_splits > 0:
            return self.fixed_num_kv_splits
        return self.lib.b200_attn_decode_num_splits(
            bs, self.num_kv_head, self.num_head // self.num_kv_head, head_dim, int(avg_len),
            self.num_sms)

    def _ensure_workspace(self, bs: int, num_splits: int) -> Optional[torch.Tensor]:
        need = self.lib.b200_attn_decode_workspace_bytes(bs, self.num_head, self.v_head_dim,
                                                         num_splits)
        if need == 0:
            return None
        if self._workspace is None or self._workspace.numel() < need:
            self._workspace = torch.empty(need, dtype=torch.uint8, device=self.device)
        return self._workspace

    def init_forward_metadata(self, forward_batch):
        bs = forward_batch.batch_size
        if forward_batch.forward_mode.is_decode_or_idle():
            if bs == 0:
                self.forward_metadata = ForwardMetadata()
                return
            total = forward_batch.seq_lens_sum or bs * self.max_context_len
            head_dim = self.v_head_dim
            s = self._num_splits(bs, head_dim, max(1, total // bs))
            self.forward_metadata = ForwardMetadata(num_splits=s,
                                                    workspace=self._ensure_workspace(bs, s))
        else:
            self.forward_metadata = self._init_extend_metadata(forward_batch)

    def _init_extend_metadata(self, forward_batch) -> ForwardMetadata:
        """qo start offsets (cumsum of extend_seq_lens) and the longest extend of the batch.
        extend_seq_lens_cpu / extend_start_loc are filled by ForwardBatch.init_new
        (forward_batch_info.py:348-369); the fallbacks mirror triton_backend.py:393-431."""
        ext = forward_batch.extend_seq_lens
        start = forward_batch.extend_start_loc
        if start is None:
            start = torch.zeros_like(ext)
            start[1:] = torch.cumsum(ext[:-1], dim=0)
        if forward_batch.extend_seq_lens_cpu is not None:
            max_len = max(forward_batch.extend_seq_lens_cpu, default=0)
            total = sum(forward_batch.extend_seq_lens_cpu)
        else:
            max_len = int(ext.max().item()) if ext.numel() else 0
            total = int(ext.sum().item()) if ext.numel() else 0
        return ForwardMetadata(qo_indptr=start.to(torch.int32), max_extend_len=int(max_len),
                               total_extend=int(total))

    # ------------------------------------------------------------------ CUDA graph hooks
    def init_cuda_graph_state(self, max_bs: int):
        self._graph_max_bs = max_bs
        self._ticket_buffer(max_bs)
        max_splits = 64
        need = self.lib.b200_attn_decode_workspace_bytes(max_bs, self.num_head, self.v_head_dim,
                                                         max_splits)
        self._graph_workspace = torch.empty(max(need, 16), dtype=torch.uint8, device=self.device)
        # int64 staging for runners that keep req_pool_indices / seq_lens as int32
        # (cuda_graph_runner.py:250-253): refreshed by the replay hook, read by the captured kernels
        self._graph_index = {n: torch.zeros(max_bs, dtype=torch.int64, device=self.device)
                             for n in ("req_pool_indices", "seq_lens")}

    def _stage_graph_indices(self, md, bs, req_pool_indices, seq_lens):
        for name, t in (("req_pool_indices", req_pool_indices), ("seq_lens", seq_lens)):
            if t is None or t.dtype == torch.int64:
                continue
            t = t[:bs]
            buf = self._graph_index[name][:bs]
            buf.copy_(t)
            if md is not None:
                md.converted[(t.data_ptr(), t.dtype, torch.int64, t.numel())] = buf

    def init_forward_metadata_capture_cuda_graph(self, bs, num_tokens, req_pool_indices, seq_lens,
                                                 encoder_lens, forward_mode, spec_info):
        if not forward_mode.is_decode_or_idle():
            raise ValueError(f"B200AttnBackend: cuda graph capture of {forward_mode!r} unsupported")
        if spec_info is not None:
            raise NotImplementedError("speculative decoding is outside the B200 hot-path scope")
        # The split count is frozen into the captured launch: choose it for a nominal context.
        s = min(64, self._num_splits(bs, self.v_head_dim, min(self.max_context_len, 4096)))
        self.forward_metadata = ForwardMetadata(num_splits=s, workspace=self._graph_workspace)
        self._stage_graph_indices(self.forward_metadata, bs, req_pool_indices, seq_lens)

    def init_forward_metadata_replay_cuda_graph(self, bs, req_pool_indices, seq_lens, seq_lens_sum,
                                                encoder_lens, forward_mode, spec_info,
                                                seq_lens_cpu):
        # The decode kernel reads req_to_token / seq_lens directly, so there is no derived index
        # buffer to rebuild and no host sync; only int32 index tensors are widened into the
        # persistent int64 staging the captured launches point at (two tiny async copies).
        self._stage_graph_indices(None, bs, req_pool_indices, seq_lens)

    def get_cuda_graph_seq_len_fill_value(self):
        return 1

    # ------------------------------------------------------------------ forward
    def _save_kv(self, layer, forward_batch, k, v):
        pool = forward_batch.token_to_kv_pool
        k_cache = pool.get_key_buffer(layer.layer_id)
        v_cache = pool.get_value_buffer(layer.layer_id)
        if k.dtype != k_cache.dtype:
            # quantised KV cache (FP8): scaling/cast order is owned by the pool
            ks, vs = self._kv_scales(layer)
            pool.set_kv_buffer(layer, forward_batch.out_cache_loc, k, v, ks, vs)
        else:
            kv_write(k, v, k_cache, v_cache, forward_batch.out_cache_loc)
        return k_cache, v_cache

    @staticmethod
    def _kv_scales(layer):
        """(k_scale, v_scale) as host floats, or (None, None) when the layer has none
        (FlashInfer passes layer.k_scale/v_scale, flashinfer_backend.py:427,499; the float copies
        are radix_attention.py:74-75)."""
        ks = getattr(layer, "k_scale_float", None)
        vs = getattr(layer, "v_scale_float", None)
        # float(tensor) is a device sync and cannot be captured in a CUDA graph: read each scale
        # once (the reference fills k_scale_float at load time, radix_attention.py:74-75) and cache
        if ks is None and layer.k_scale is not None:
            ks = layer.k_scale_float = float(layer.k_scale)
        if vs is None and layer.v_scale is not None:
            vs = layer.v_scale_float = float(layer.v_scale)
        return ks, vs

    @staticmethod
    def _check_layer(layer):
        """Features of the reference's Triton backend that this backend does not implement must
        fail loudly instead of silently running full self-attention (triton_backend.py:271-446
        handles the window; cross-attention indexes encoder rows)."""
        window = getattr(layer, "sliding_window_size", None)
        if window is not None and window > 0:
            raise NotImplementedError(
                f"B200AttnBackend: sliding_window_size={window} is not implemented "
                "(no BASELINE config uses a window)")
        if getattr(layer, "is_cross_attention", False):
            raise NotImplementedError("B200AttnBackend: cross-attention layers are not implemented")

    def forward_decode(self, q, k, v, layer, forward_batch, save_kv_cache: bool = True,
                       fp8_out: bool = False):
        """fp8_out (B200 extension, default off = the reference contract): also quantise the output
        row per token to e4m3 inside the attention kernel and return a QuantizedActivation for a
        per-token FP8 o_proj, instead of the 16-bit tensor + a separate quantiser launch."""
        hq, hkv = layer.tp_q_head_num, layer.tp_k_head_num
        d, dv = layer.qk_head_dim, layer.v_head_dim
        q = q.reshape(-1, hq * d)
        if q.stride(-1) != 1:
            q = q.contiguous()
        bs = q.shape[0]
        o = q.new_empty((bs, hq * dv))
        pool = forward_batch.token_to_kv_pool
        if save_kv_cache:
            k_cache, v_cache = self._save_kv(layer, forward_batch, k, v)
        else:
            k_cache = pool.get_key_buffer(layer.layer_id)
            v_cache = pool.get_value_buffer(layer.layer_id)
        if bs == 0:
            return self._maybe_quant(o, fp8_out)
        self._check_layer(layer)
        md = self.forward_metadata
        req_to_token = forward_batch.req_to_token_pool.req_to_token
        ws = md.workspace
        # the reference's CudaGraphRunner keeps these as int32 (cuda_graph_runner.py:250-253); the
        # kernels read int64, so convert once per forward and hold the copies in the metadata
        req_pool_indices = md.as_dtype(forward_batch.req_pool_indices, torch.int64)
        seq_lens = md.as_dtype(forward_batch.seq_lens, torch.int64)
        fp8kv = _lib.is_fp8_kv(k_cache.dtype)
        if not fp8kv and k_cache.dtype != q.dtype:
            raise _lib.B200Error(
                f"B200AttnBackend.forward_decode: KV cache dtype {k_cache.dtype} != q dtype "
                f"{q.dtype} (supported pools: the activation dtype, float8_e4m3fn or float8_e5m2)")
        common = (q.data_ptr(), q.stride(0), k_cache.data_ptr(), v_cache.data_ptr(),
                  k_cache.stride(0), o.data_ptr(), o.stride(0), req_to_token.data_ptr(),
                  req_to_token.stride(0), req_pool_indices.data_ptr(), seq_lens.data_ptr(), bs, hq,
                  hkv, d, dv, float(layer.scaling), float(layer.logit_cap or 0.0))
        if fp8kv:
            ks, vs = self._kv_scales(layer)
            common = common + (1.0 if ks is None else ks, 1.0 if vs is None else vs)
        common = common + (_lib.kv_dtype_code(q.dtype, k_cache.dtype), _lib.ptr(ws),
                           0 if ws is None else ws.numel(), md.num_splits)
        # the tail only pays when it replaces a launch: the merge (splits > 1) or the quantiser
        fused = self.fused_tail and (md.num_splits > 1 or fp8_out) and (hq * dv) % 8 == 0
        if fused:
            qa = sa = None
            if fp8_out:
                qa = torch.empty((bs, hq * dv), dtype=torch.float8_e4m3fn, device=q.device)
                sa = torch.empty((bs, 1), dtype=torch.float32, device=q.device)
            fn = self.lib.b200_attn_decode_fp8kv_fused if fp8kv else self.lib.b200_attn_decode_fused
            tickets = self._ticket_buffer(bs) if bs <= self.tail_in_kernel_max_bs else None
            rc = fn(*common, _lib.ptr(tickets), 0 if tickets is None else tickets.numel() * 4,
                    _lib.ptr(qa), 0 if qa is None else qa.stride(0), _lib.ptr(sa),
                    _lib.current_stream())
            _lib.check(rc, "b200_attn_decode_fused")
            if fp8_out:
                from ..quantization.fp8_utils import QuantizedActivation

                return QuantizedActivation(qa, sa, q.dtype)
            return o
        fn = self.lib.b200_attn_decode_fp8kv if fp8kv else self.lib.b200_attn_decode
        rc = fn(*common, _lib.current_stream())
        _lib.check(rc, "b200_attn_decode")
        return self._maybe_quant(o, fp8_out)

    @staticmethod
    def _maybe_quant(o, fp8_out):
        if not fp8_out:
            return o
        from ..quantization.fp8_kernel import sglang_per_token_quant_fp8 as per_token_quant_fp8
        from ..quantization.fp8_utils import QuantizedActivation

        qa, sa = per_token_quant_fp8(o)
        return QuantizedActivation(qa, sa, o.dtype)

    def forward_extend(self, q, k, v, layer, forward_batch, save_kv_cache: bool = True, **_):
        hq, hkv = layer.tp_q_head_num, layer.tp_k_head_num
        d, dv = layer.qk_head_dim, layer.v_head_dim
        q = q.reshape(-1, hq * d)
        if q.stride(-1) != 1:
            q = q.contiguous()
        o = q.new_empty((q.shape[0], hq * dv))
        pool = forward_batch.token_to_kv_pool
        if save_kv_cache:
            # new tokens' K/V go to the pool first; the kernel then gathers prefix AND new keys
            # from the pool, exactly like the oracle (torch_native_backend.py:196-199)
            k_cache, v_cache = self._save_kv(layer, forward_batch, k, v)
        else:
            k_cache = pool.get_key_buffer(layer.layer_id)
            v_cache = pool.get_value_buffer(layer.layer_id)
        md = self.forward_metadata
        bs = forward_batch.batch_size
        if q.shape[0] == 0 or bs == 0 or md.max_extend_len == 0:
            return o
        attn_type = getattr(layer, "attn_type", None)
        causal = not (layer.is_cross_attention or
                      getattr(attn_type, "value", attn_type) == "encoder_only")
        req_to_token = forward_batch.req_to_token_pool.req_to_token
        self._check_layer(layer)
        req_pool_indices = md.as_dtype(forward_batch.req_pool_indices, torch.int64)
        seq_lens = md.as_dtype(forward_batch.seq_lens, torch.int64)
        prefix_lens = md.as_dtype(forward_batch.extend_prefix_lens, torch.int32)
        extend_lens = md.as_dtype(forward_batch.extend_seq_lens, torch.int32)
        scaling, out_scale = float(layer.scaling), 1.0
        if _lib.is_fp8_kv(k_cache.dtype):
            # FP8 pool: the tcgen05 kernel wants 16-bit operands, so the requests' rows are first
            # copied (exactly) into a dense 16-bit scratch; scales ride on the softmax scale / output
            k_cache, v_cache, req_to_token, req_pool_indices = self._dequant_for_extend(
                forward_batch, k_cache, v_cache, req_to_token, req_pool_indices, seq_lens, hkv * d,
                q.dtype)
            ks, vs = self._kv_scales(layer)
            scaling *= 1.0 if ks is None else ks
            out_scale = 1.0 if vs is None else vs
        elif k_cache.dtype != q.dtype:
            raise _lib.B200Error(
                f"B200AttnBackend.forward_extend: KV cache dtype {k_cache.dtype} != q dtype "
                f"{q.dtype} (supported pools: the activation dtype, float8_e4m3fn or float8_e5m2)")
        rc = self.lib.b200_attn_extend_scaled(
            q.data_ptr(), q.stride(0), k_cache.data_ptr(), v_cache.data_ptr(), k_cache.stride(0),
            o.data_ptr(), o.stride(0), req_to_token.data_ptr(), req_to_token.stride(0),
            req_pool_indices.data_ptr(), seq_lens.data_ptr(),
            prefix_lens.data_ptr(), extend_lens.data_ptr(), md.qo_indptr.data_ptr(), bs,
            md.max_extend_len, hq, hkv, d, dv, scaling, float(layer.logit_cap or 0.0), out_scale,
            1 if causal else 0, _lib.dtype_code(q.dtype), _lib.current_stream())
        _lib.check(rc, "b200_attn_extend")
        return o

    def _dequant_for_extend(self, forward_batch, k_cache, v_cache, req_to_token, req_pool_indices,
                            seq_lens, row_elems, dtype):
        """-> (k16, v16, dense_req_to_token, dense_req_pool_indices) for an e4m3 / e5m2 pool."""
        bs = seq_lens.numel()
        cpu = getattr(forward_batch, "seq_lens_cpu", None)
        if cpu is not None:
            total, max_len = int(cpu.sum()), int(cpu.max())
        else:
            total, max_len = int(seq_lens.sum().item()), int(seq_lens.max().item())
        indptr = torch.zeros(bs + 1, dtype=torch.int32, device=seq_lens.device)
        indptr[1:] = torch.cumsum(seq_lens, dim=0)
        k16 = torch.empty((total, row_elems), dtype=dtype, device=seq_lens.device)
        v16 = torch.empty_like(k16)
        stride = (max_len + 3) // 4 * 4
        dense_rtt = torch.empty((bs, stride), dtype=torch.int32, device=seq_lens.device)
        rc = self.lib.b200_kv_dequant_gather(
            k_cache.data_ptr(), v_cache.data_ptr(), row_elems, req_to_token.data_ptr(),
            req_to_token.stride(0), req_pool_indices.data_ptr(), seq_lens.data_ptr(),
            indptr.data_ptr(), k16.data_ptr(), v16.data_ptr(), dense_rtt.data_ptr(), stride, bs,
            max_len, _lib.kv_dtype_code(dtype, k_cache.dtype), _lib.current_stream())
        _lib.check(rc, "b200_kv_dequant_gather")
        return k16, v16, dense_rtt, torch.arange(bs, dtype=torch.int64, device=seq_lens.device)

    def support_triton(self):
        return False
