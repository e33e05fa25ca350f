"""Linear layers with the reference's structure (python/sglang/srt/layers/linear.py): the quant
method owns the weights and the math (`quant_method.apply(layer, x, bias)`), column-parallel layers
shard N, row-parallel layers shard K and all-reduce their partial outputs (linear.py:1277-1299;
bias only on rank 0, :1290).  Weight loaders slice checkpoints by tp_rank (linear.py:352-358, 1184).
"""
from __future__ import annotations

from typing import List, Optional

import torch
import torch.nn.functional as F
from torch.nn import Parameter

from ..distributed.parallel_state import (get_tensor_model_parallel_rank,
                                          get_tensor_model_parallel_world_size,
                                          get_tp_custom_all_reduce,
                                          tensor_model_parallel_all_gather,
                                          tensor_model_parallel_all_reduce)
from .dense_kernel import dense_supported, gemm_16bit
from .quantization.base_config import LinearMethodBase, QuantizationConfig, set_weight_attrs


def divide(a: int, b: int) -> int:
    if a % b != 0:
        raise ValueError(f"{a} is not divisible by {b}")
    return a // b


class UnquantizedLinearMethod(LinearMethodBase):
    """16-bit weights (linear.py:143-176).  CUDA bf16/f16 operands run on the in-tree tcgen05 dense
    GEMM (csrc/dense_gemm.cu); anything else (the CPU tensors of the gloo host-logic tests, fp32)
    is torch's F.linear exactly as in the reference."""

    def create_weights(self, layer, input_size_per_partition, output_partition_sizes, input_size,
                       output_size, params_dtype, **extra_weight_attrs):
        extra_weight_attrs.pop("tp_size", None)
        weight = Parameter(torch.empty(sum(output_partition_sizes), input_size_per_partition,
                                       dtype=params_dtype), requires_grad=False)
        set_weight_attrs(weight, {"input_dim": 1, "output_dim": 0})
        layer.register_parameter("weight", weight)
        set_weight_attrs(weight, extra_weight_attrs)

    def apply(self, layer, x, bias=None):
        if dense_supported(x, layer.weight):
            return gemm_16bit(x, layer.weight, bias)
        return F.linear(x, layer.weight, bias)


class LinearBase(torch.nn.Module):
    def __init__(self, input_size: int, output_size: int, skip_bias_add: bool = False,
                 params_dtype: Optional[torch.dtype] = None,
                 quant_config: Optional[QuantizationConfig] = None, prefix: str = ""):
        super().__init__()
        self.input_size = input_size
        self.output_size = output_size
        self.skip_bias_add = skip_bias_add
        self.params_dtype = params_dtype or torch.get_default_dtype()
        self.quant_method = (UnquantizedLinearMethod() if quant_config is None
                             else quant_config.get_quant_method(self, prefix=prefix))
        self.prefix = prefix

    def forward(self, x):
        raise NotImplementedError()


class ColumnParallelLinear(LinearBase):
    def __init__(self, input_size: int, output_size: int, bias: bool = True,
                 gather_output: bool = False, skip_bias_add: bool = False,
                 params_dtype: Optional[torch.dtype] = None,
                 quant_config: Optional[QuantizationConfig] = None,
                 output_sizes: Optional[List[int]] = None, prefix: str = "",
                 tp_rank: Optional[int] = None, tp_size: Optional[int] = None):
        super().__init__(input_size, output_size, skip_bias_add, params_dtype, quant_config, prefix)
        self.gather_output = gather_output
        self.tp_rank = get_tensor_model_parallel_rank() if tp_rank is None else tp_rank
        self.tp_size = get_tensor_model_parallel_world_size() if tp_size is None else tp_size
        self.output_size_per_partition = divide(output_size, self.tp_size)
        self.output_partition_sizes = [self.output_size_per_partition]
        if output_sizes is not None:  # fused layers (qkv / gate_up): one logical width per shard
            self.output_partition_sizes = [divide(s, self.tp_size) for s in output_sizes]
        self.quant_method.create_weights(
            layer=self, input_size_per_partition=input_size,
            output_partition_sizes=self.output_partition_sizes, input_size=input_size,
            output_size=output_size, params_dtype=self.params_dtype,
            weight_loader=self.weight_loader, tp_size=self.tp_size)
        if bias:
            self.bias = Parameter(torch.zeros(self.output_size_per_partition, dtype=self.params_dtype),
                                  requires_grad=False)
            set_weight_attrs(self.bias, {"output_dim": 0, "weight_loader": self.weight_loader})
        else:
            self.register_parameter("bias", None)

    def weight_loader(self, param: Parameter, loaded_weight: torch.Tensor):
        output_dim = getattr(param, "output_dim", None)
        if output_dim is not None and loaded_weight.shape[output_dim] != param.shape[output_dim]:
            shard = param.shape[output_dim]
            loaded_weight = loaded_weight.narrow(output_dim, self.tp_rank * shard, shard)
        if loaded_weight.dim() == 0:
            loaded_weight = loaded_weight.reshape(1)
        assert param.shape == loaded_weight.shape, (param.shape, loaded_weight.shape)
        param.data.copy_(loaded_weight)

    # ---- fused layers (gate_up, qkv): the reference's `loaded_shard_id` contract ----------------
    def _load_logical_shard(self, param: Parameter, loaded_weight: torch.Tensor, slot: int,
                            offset: int, size: int, rank_index: int):
        """Copy one logical projection of a fused parameter.  `offset`/`size` are this rank's
        window in UNPACKED output rows (linear.py:615-640 / 1060-1110); `rank_index` selects the
        slice of the checkpoint tensor (tp_rank, or tp_rank // replicas for replicated kv heads).
        int4-packed and block-scale parameters rescale the window
        (parameter.py:340-375 `adjust_shard_indexes_for_packing`, linear.py:731-741)."""
        output_dim = getattr(param, "output_dim", None)
        if output_dim is None:
            # one scalar per logical shard (Fp8 per-tensor scales, linear.py:78-94
            # `adjust_scalar_to_fused_array`): slot = logical shard index
            if loaded_weight.numel() != 1:
                raise ValueError(f"fused parameter without output_dim expects a scalar per shard, "
                                 f"got {tuple(loaded_weight.shape)}")
            param.data[slot].copy_(loaded_weight.reshape(()))
            return
        if getattr(param, "packed_dim", None) == output_dim:
            factor = param.packed_factor
            if offset % factor or size % factor:
                raise ValueError(f"shard window ({offset}, {size}) is not a multiple of the pack "
                                 f"factor {factor}")
            offset, size = offset // factor, size // factor
        block_n = getattr(param, "block_n", None)
        if block_n is not None:  # [N/block_n, K/block_k] scale grid
            offset, size = (offset + block_n - 1) // block_n, (size + block_n - 1) // block_n
        window = param.data.narrow(output_dim, offset, size)
        if loaded_weight.shape[output_dim] != size:
            if loaded_weight.shape[output_dim] < (rank_index + 1) * size:
                raise ValueError(f"checkpoint shard of width {loaded_weight.shape[output_dim]} has no "
                                 f"window {rank_index} of size {size}")
            loaded_weight = loaded_weight.narrow(output_dim, rank_index * size, size)
        if window.shape != loaded_weight.shape:
            raise ValueError(f"shard shape {tuple(loaded_weight.shape)} does not fit the "
                             f"parameter window {tuple(window.shape)}")
        window.copy_(loaded_weight)

    def _split_fused_checkpoint(self, param: Parameter, loaded_weight: torch.Tensor, ids, sizes):
        """A tensor already fused on disk (Phi-3 style, linear.py:560-606, 983-1040): cut it into
        its logical projections (FULL, un-sharded widths) and load each with its shard id."""
        output_dim = getattr(param, "output_dim", None)
        if output_dim is None:
            if loaded_weight.numel() == 1:      # one per-tensor scale for the fused tensor
                param.data.fill_(loaded_weight.reshape(()).item())
            else:
                assert param.shape == loaded_weight.shape, (param.shape, loaded_weight.shape)
                param.data.copy_(loaded_weight)
            return
        start = 0
        for sid, full in zip(ids, sizes):
            lo, width = start, full
            if getattr(param, "packed_dim", None) == output_dim:
                lo, width = lo // param.packed_factor, width // param.packed_factor
            block_n = getattr(param, "block_n", None)
            if block_n is not None:
                lo, width = (lo + block_n - 1) // block_n, (width + block_n - 1) // block_n
            self.weight_loader(param, loaded_weight.narrow(output_dim, lo, width), sid)
            start += full

    def forward(self, input_):
        bias = self.bias if not self.skip_bias_add else None
        output = self.quant_method.apply(self, input_, bias)
        if self.gather_output:
            output = tensor_model_parallel_all_gather(output)
        return output, (self.bias if self.skip_bias_add else None)


class MergedColumnParallelLinear(ColumnParallelLinear):
    """Several column-parallel projections fused along N (gate_up_proj)."""

    def __init__(self, input_size: int, output_sizes: List[int], bias: bool = True, **kw):
        self.output_sizes = output_sizes
        super().__init__(input_size, sum(output_sizes), bias=bias, output_sizes=output_sizes, **kw)

    def weight_loader(self, param: Parameter, loaded_weight: torch.Tensor,
                      loaded_shard_id: Optional[int] = None):
        """linear.py:523-660: `loaded_shard_id` = index of the logical projection (0 = gate,
        1 = up); None = the checkpoint tensor is already fused."""
        if loaded_shard_id is None:
            self._split_fused_checkpoint(param, loaded_weight, range(len(self.output_sizes)),
                                         self.output_sizes)
            return
        if not 0 <= loaded_shard_id < len(self.output_sizes):
            raise ValueError(f"loaded_shard_id {loaded_shard_id} out of range")
        offset = sum(self.output_sizes[:loaded_shard_id]) // self.tp_size
        size = self.output_sizes[loaded_shard_id] // self.tp_size
        self._load_logical_shard(param, loaded_weight, loaded_shard_id, offset, size, self.tp_rank)


class QKVParallelLinear(ColumnParallelLinear):
    """Fused q/k/v projection; q heads / tp, kv heads max(1, Hkv // tp) (replicated when tp > Hkv),
    as models/llama.py:119-132."""

    def __init__(self, hidden_size: int, head_size: int, total_num_heads: int,
                 total_num_kv_heads: Optional[int] = None, bias: bool = True, **kw):
        tp_size = kw.get("tp_size") or get_tensor_model_parallel_world_size()
        self.hidden_size, self.head_size = hidden_size, head_size
        self.total_num_heads = total_num_heads
        self.total_num_kv_heads = total_num_kv_heads or total_num_heads
        self.num_heads = divide(total_num_heads, tp_size)
        if tp_size >= self.total_num_kv_heads:
            self.num_kv_heads, self.num_kv_head_replicas = 1, divide(tp_size, self.total_num_kv_heads)
        else:
            self.num_kv_heads, self.num_kv_head_replicas = divide(self.total_num_kv_heads, tp_size), 1
        out = (self.num_heads + 2 * self.num_kv_heads) * tp_size * head_size
        sizes = [self.num_heads * head_size * tp_size, self.num_kv_heads * head_size * tp_size,
                 self.num_kv_heads * head_size * tp_size]
        super().__init__(hidden_size, out, bias=bias, output_sizes=sizes, **kw)

    _SLOT = {"q": 0, "k": 1, "v": 2}

    def weight_loader(self, param: Parameter, loaded_weight: torch.Tensor,
                      loaded_shard_id: Optional[str] = None):
        """linear.py:937-1128: `loaded_shard_id` in {"q", "k", "v"}; kv heads replicated over
        `num_kv_head_replicas` ranks read the same checkpoint slice (tp_rank // replicas)."""
        hd = self.head_size
        if loaded_shard_id is None:
            self._split_fused_checkpoint(
                param, loaded_weight, ("q", "k", "v"),
                (self.total_num_heads * hd, self.total_num_kv_heads * hd,
                 self.total_num_kv_heads * hd))
            return
        if loaded_shard_id not in self._SLOT:
            raise ValueError(f"loaded_shard_id must be 'q', 'k' or 'v', got {loaded_shard_id!r}")
        q, kv = self.num_heads * hd, self.num_kv_heads * hd
        offset, size = {"q": (0, q), "k": (q, kv), "v": (q + kv, kv)}[loaded_shard_id]
        rank_index = (self.tp_rank if loaded_shard_id == "q"
                      else self.tp_rank // self.num_kv_head_replicas)
        self._load_logical_shard(param, loaded_weight, self._SLOT[loaded_shard_id], offset, size,
                                 rank_index)


class RowParallelLinear(LinearBase):
    def __init__(self, input_size: int, output_size: int, bias: bool = True,
                 input_is_parallel: bool = True, skip_bias_add: bool = False,
                 params_dtype: Optional[torch.dtype] = None, reduce_results: bool = True,
                 quant_config: Optional[QuantizationConfig] = None, prefix: str = "",
                 tp_rank: Optional[int] = None, tp_size: Optional[int] = None):
        super().__init__(input_size, output_size, skip_bias_add, params_dtype, quant_config, prefix)
        self.input_is_parallel = input_is_parallel
        self.reduce_results = reduce_results
        self.tp_rank = get_tensor_model_parallel_rank() if tp_rank is None else tp_rank
        self.tp_size = get_tensor_model_parallel_world_size() if tp_size is None else tp_size
        self.input_size_per_partition = divide(input_size, self.tp_size)
        self.quant_method.create_weights(
            layer=self, input_size_per_partition=self.input_size_per_partition,
            output_partition_sizes=[output_size], input_size=input_size, output_size=output_size,
            params_dtype=self.params_dtype, weight_loader=self.weight_loader, tp_size=self.tp_size)
        if not reduce_results and bias and not skip_bias_add:
            raise ValueError("When not reduce the results, adding bias to the results can lead to "
                             "incorrect results")
        if bias:
            self.bias = Parameter(torch.zeros(output_size, dtype=self.params_dtype),
                                  requires_grad=False)
        else:
            self.register_parameter("bias", None)

    def weight_loader(self, param: Parameter, loaded_weight: torch.Tensor):
        input_dim = getattr(param, "input_dim", None)
        if input_dim is not None and loaded_weight.shape[input_dim] != param.shape[input_dim]:
            shard = param.shape[input_dim]
            loaded_weight = loaded_weight.narrow(input_dim, self.tp_rank * shard, shard)
        if loaded_weight.dim() == 0:
            loaded_weight = loaded_weight.reshape(1)
        assert param.shape == loaded_weight.shape, (param.shape, loaded_weight.shape)
        param.data.copy_(loaded_weight)

    def forward(self, input_):
        if not self.input_is_parallel:
            input_ = input_.chunk(self.tp_size, dim=-1)[self.tp_rank].contiguous()
        # bias is fused into the GEMM only on rank 0 so the all-reduce adds it exactly once
        bias_ = None if (self.tp_rank > 0 or self.skip_bias_add) else self.bias
        if self.reduce_results and self.tp_size > 1:
            comm = get_tp_custom_all_reduce()
            # the peer-memory path (fused epilogue or stand-alone one-shot kernel, chosen inside the
            # library) only up to the measured crossover with NCCL for this group size
            # (input_ may be a QuantizedActivation: only .shape and .dtype = the output dtype)
            tokens = 1
            for dim in input_.shape[:-1]:
                tokens *= int(dim)
            out_bytes = tokens * self.output_size * input_.dtype.itemsize
            if (comm is not None and out_bytes <= getattr(comm, "crossover", 0)
                    and getattr(self.quant_method, "supports_fused_all_reduce", False)):
                # GEMM -> all-reduce as one call: the decode-regime kernel reduces in its epilogue
                self.fused_all_reduce = comm
                try:
                    output = self.quant_method.apply(self, input_, bias=bias_)
                finally:
                    self.fused_all_reduce = None
            else:
                output = self.quant_method.apply(self, input_, bias=bias_)
                output = tensor_model_parallel_all_reduce(output)
        else:
            output = self.quant_method.apply(self, input_, bias=bias_)
        return output, (self.bias if self.skip_bias_add else None)
