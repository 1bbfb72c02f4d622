"""Tensor-parallel linear wrappers around a quantized linear method.

Only what the PQ path needs from vllm/model_executor/layers/linear.py:
  ReplicatedLinear            :197-283    ColumnParallelLinear        :285-443
  MergedColumnParallelLinear  :445-717    QKVParallelLinear           :719-1058
  RowParallelLinear           :1061-1230
Same constructor arguments, attribute names (output_partition_sizes,
input_size_per_partition, ...) and ``forward`` contract; the weight loaders
implement the legacy attribute-driven protocol (output_dim / input_dim /
packed_dim / pack_factor / is_metadata) for the three PQ
parameters and a bias.  Sharding rules:
  column-parallel: rows of codes/scales narrowed by rank; codebooks replicated,
                   fused shards place their M codebooks at shard_index*M
  row-parallel:    codes narrowed on the group axis; scales, codebooks whole;
                   bias on rank 0 only; one all-reduce after apply
"""
from __future__ import annotations

from typing import List, Optional, Tuple, Union

import torch
from torch import nn
from torch.nn.parameter import Parameter

from .distributed import (get_tensor_model_parallel_rank, get_tensor_model_parallel_world_size,
                          tensor_model_parallel_all_gather, tensor_model_parallel_all_reduce)
from .plugin_api import QuantizationConfig, set_weight_attrs


def divide(numerator: int, denominator: int) -> int:
    assert numerator % denominator == 0, f"{numerator} is not divisible by {denominator}"
    return numerator // denominator


class LinearBase(nn.Module):
    """linear.py:146-194"""

    def __init__(self, input_size: int, output_size: int, skip_bias_add: bool = False,
                 params_dtype: Optional[torch.dtype] = None,
                 quant_config: Optional[QuantizationConfig] = None, prefix: str = "",
                 *, return_bias: bool = True):
        super().__init__()
        self.input_size = input_size
        self.output_size = output_size
        self.skip_bias_add = skip_bias_add
        self.params_dtype = params_dtype if params_dtype is not None else torch.get_default_dtype()
        if quant_config is None:
            raise ValueError("vllm_pq_b200 layers need a quantization config (PQConfig)")
        self.quant_method = quant_config.get_quant_method(self, prefix=prefix)
        self.return_bias = return_bias

    def _ret(self, output, output_bias):
        return (output, output_bias) if self.return_bias else output


def is_linear_layer(layer: nn.Module) -> bool:
    if isinstance(layer, LinearBase):
        return True
    try:  # a real vLLM layer, when this method is plugged into vLLM
        import sys
        mod = sys.modules.get("vllm.model_executor.layers.linear")
        return mod is not None and isinstance(layer, mod.LinearBase)
    except Exception:
        return False


def _copy(param_data: torch.Tensor, loaded: torch.Tensor):
    assert param_data.shape == loaded.shape, (tuple(param_data.shape), tuple(loaded.shape))
    param_data.copy_(loaded)


class ReplicatedLinear(LinearBase):
    def __init__(self, input_size, output_size, bias=True, skip_bias_add=False, params_dtype=None,
                 quant_config=None, prefix="", *, return_bias=True):
        super().__init__(input_size, output_size, skip_bias_add, params_dtype, quant_config, prefix,
                         return_bias=return_bias)
        self.quant_method.create_weights(self, self.input_size, [self.output_size], self.input_size,
                                         self.output_size, self.params_dtype,
                                         weight_loader=self.weight_loader)
        if bias:
            self.bias = Parameter(torch.empty(self.output_size, dtype=self.params_dtype))
            set_weight_attrs(self.bias, {"output_dim": 0, "weight_loader": self.weight_loader})
        else:
            self.register_parameter("bias", None)

    def weight_loader(self, param: Parameter, loaded_weight: torch.Tensor):
        _copy(param.data, loaded_weight)

    def forward(self, x):
        bias = self.bias if not self.skip_bias_add else None
        output = self.quant_method.apply(self, x, bias)
        return self._ret(output, self.bias if self.skip_bias_add else None)


class ColumnParallelLinear(LinearBase):
    def __init__(self, input_size, output_size, bias=True, gather_output=False, skip_bias_add=False,
                 params_dtype=None, quant_config=None, output_sizes: Optional[List[int]] = None,
                 prefix="", *, return_bias=True):
        self.tp_size = get_tensor_model_parallel_world_size()
        self.input_size_per_partition = input_size
        self.output_size_per_partition = divide(output_size, self.tp_size)
        self.output_partition_sizes = [self.output_size_per_partition]
        if hasattr(self, "output_sizes"):            # fused layers: one partition per logical matrix
            self.output_partition_sizes = [divide(s, self.tp_size) for s in self.output_sizes]
        super().__init__(input_size, output_size, skip_bias_add, params_dtype, quant_config, prefix,
                         return_bias=return_bias)
        self.gather_output = gather_output
        self.quant_method.create_weights(
            layer=self, input_size_per_partition=self.input_size_per_partition,
            output_partition_sizes=self.output_partition_sizes, input_size=self.input_size,
            output_size=self.output_size, params_dtype=self.params_dtype,
            weight_loader=self.weight_loader)
        if bias:
            self.bias = Parameter(torch.empty(self.output_size_per_partition, dtype=self.params_dtype))
            set_weight_attrs(self.bias, {"output_dim": 0, "weight_loader": self.weight_loader})
        else:
            self.register_parameter("bias", None)

    def weight_loader(self, param: Parameter, loaded_weight: torch.Tensor):
        output_dim = getattr(param, "output_dim", None)
        if output_dim is not None:
            shard = param.data.shape[output_dim]
            loaded_weight = loaded_weight.narrow(output_dim, get_tensor_model_parallel_rank() * shard, shard)
        _copy(param.data, loaded_weight)

    def forward(self, input_, input_norm=None) -> Union[torch.Tensor, Tuple[torch.Tensor, Optional[Parameter]]]:
        """`input_norm=(ssq, weight, eps)`: `input_` is the residual stream and the layer's input is
        RMSNorm(input_) * weight (llama.py:272-283) - taken on the kernel's x load at decode, with the
        sum of squares `ssq` handed over by the row-parallel layer that produced `input_`
        (`want_ssq=True`; None: computed by a norm launch).  SURVEY.md 8f, N1."""
        bias = self.bias if not self.skip_bias_add else None
        if input_norm is not None:
            output, _ = self.quant_method.apply_fused(self, input_, bias, norm=input_norm)
        else:
            output = self.quant_method.apply(self, input_, bias)      # no collective (linear.py:425)
        if self.gather_output:
            output = tensor_model_parallel_all_gather(output)
        return self._ret(output, self.bias if self.skip_bias_add else None)


class MergedColumnParallelLinear(ColumnParallelLinear):
    """gate_up_proj: several column-parallel matrices concatenated on the output axis."""

    def __init__(self, input_size, output_sizes: List[int], bias=True, gather_output=False,
                 skip_bias_add=False, params_dtype=None, quant_config=None, prefix="", *,
                 return_bias=True):
        self.output_sizes = output_sizes
        tp = get_tensor_model_parallel_world_size()
        assert all(s % tp == 0 for s in output_sizes)
        super().__init__(input_size, sum(output_sizes), bias, gather_output, skip_bias_add,
                         params_dtype, quant_config, prefix=prefix, return_bias=return_bias)

    def _shards(self):
        off = 0
        for i, s in enumerate(self.output_sizes):
            yield i, off, s
            off += s

    def weight_loader(self, param: Parameter, loaded_weight: torch.Tensor,
                      loaded_shard_id: Optional[int] = None):
        output_dim = getattr(param, "output_dim", None)
        is_metadata = getattr(param, "is_metadata", False)
        if loaded_shard_id is None:                  # checkpoint already fused on disk
            if output_dim is None:
                _copy(param.data, loaded_weight)
                return
            packed = getattr(param, "packed_dim", None) == output_dim
            for i, off, size in self._shards():
                if packed:
                    off, size = off // param.pack_factor, size // param.pack_factor
                self.weight_loader(param, loaded_weight.narrow(output_dim, off, size), i)
            return
        assert loaded_shard_id < len(self.output_sizes)
        tp_rank, tp_size = get_tensor_model_parallel_rank(), get_tensor_model_parallel_world_size()
        data = param.data
        if output_dim is not None:
            off = sum(self.output_sizes[:loaded_shard_id]) // tp_size
            size = self.output_sizes[loaded_shard_id] // tp_size
            if getattr(param, "packed_dim", None) == output_dim:
                off, size = off // param.pack_factor, size // param.pack_factor
            data = data.narrow(output_dim, off, size)
            loaded_weight = loaded_weight.narrow(output_dim, tp_rank * size, size)
        elif is_metadata:                            # codebooks: M books per shard, replicated
            size = loaded_weight.shape[0]
            data = data.narrow(0, loaded_shard_id * size, size)
        _copy(data, loaded_weight)


class QKVParallelLinear(ColumnParallelLinear):
    """qkv_proj: q heads split by rank, kv heads split or replicated (linear.py:759-792)."""

    def __init__(self, hidden_size, head_size, total_num_heads, total_num_kv_heads=None, bias=True,
                 skip_bias_add=False, params_dtype=None, quant_config=None, prefix="", *,
                 return_bias=True):
        self.hidden_size, self.head_size = hidden_size, head_size
        self.total_num_heads = total_num_heads
        self.total_num_kv_heads = total_num_kv_heads if total_num_kv_heads is not None else total_num_heads
        tp = get_tensor_model_parallel_world_size()
        self.num_heads = divide(self.total_num_heads, tp)
        if tp >= self.total_num_kv_heads:
            self.num_kv_heads, self.num_kv_head_replicas = 1, divide(tp, self.total_num_kv_heads)
        else:
            self.num_kv_heads, self.num_kv_head_replicas = divide(self.total_num_kv_heads, tp), 1
        output_size = (self.num_heads + 2 * self.num_kv_heads) * tp * self.head_size
        self.output_sizes = [self.num_heads * self.head_size * tp,
                             self.num_kv_heads * self.head_size * tp,
                             self.num_kv_heads * self.head_size * tp]
        super().__init__(hidden_size, output_size, bias, False, skip_bias_add, params_dtype,
                         quant_config, prefix=prefix, return_bias=return_bias)

    def _shard_span(self, shard_id: str) -> Tuple[int, int]:
        q = self.num_heads * self.head_size
        kv = self.num_kv_heads * self.head_size
        return {"q": (0, q), "k": (q, kv), "v": (q + kv, kv)}[shard_id]

    def weight_loader(self, param: Parameter, loaded_weight: torch.Tensor,
                      loaded_shard_id: Optional[str] = None):
        output_dim = getattr(param, "output_dim", None)
        is_metadata = getattr(param, "is_metadata", False)
        if loaded_shard_id is None:                  # fused qkv on disk
            if output_dim is None:
                _copy(param.data, loaded_weight)
                return
            tq = self.total_num_heads * self.head_size
            tkv = self.total_num_kv_heads * self.head_size
            packed = getattr(param, "packed_dim", None) == output_dim
            for sid, off, size in (("q", 0, tq), ("k", tq, tkv), ("v", tq + tkv, tkv)):
                if packed:
                    off, size = off // param.pack_factor, size // param.pack_factor
                self.weight_loader(param, loaded_weight.narrow(output_dim, off, size), sid)
            return
        assert loaded_shard_id in ("q", "k", "v")
        tp_rank = get_tensor_model_parallel_rank()
        data = param.data
        if output_dim is not None:
            off, size = self._shard_span(loaded_shard_id)
            if getattr(param, "packed_dim", None) == output_dim:
                off, size = off // param.pack_factor, size // param.pack_factor
            data = data.narrow(output_dim, off, size)
            shard = tp_rank if loaded_shard_id == "q" else tp_rank // self.num_kv_head_replicas
            loaded_weight = loaded_weight.narrow(output_dim, shard * size, size)
        elif is_metadata:
            size = loaded_weight.shape[0]
            data = data.narrow(0, ["q", "k", "v"].index(loaded_shard_id) * size, size)
        _copy(data, loaded_weight)


class RowParallelLinear(LinearBase):
    """o_proj / down_proj: K split by rank, partial sums all-reduced (linear.py:1193-1221)."""

    def __init__(self, input_size, output_size, bias=True, input_is_parallel=True,
                 skip_bias_add=False, params_dtype=None, reduce_results=True, quant_config=None,
                 prefix="", *, return_bias=True):
        self.tp_rank = get_tensor_model_parallel_rank()
        self.tp_size = get_tensor_model_parallel_world_size()
        self.input_size_per_partition = divide(input_size, self.tp_size)
        self.output_size_per_partition = output_size
        self.output_partition_sizes = [output_size]
        super().__init__(input_size, output_size, skip_bias_add, params_dtype, quant_config, prefix,
                         return_bias=return_bias)
        self.input_is_parallel = input_is_parallel
        self.reduce_results = reduce_results
        self.quant_method.create_weights(
            layer=self, input_size_per_partition=self.input_size_per_partition,
            output_partition_sizes=self.output_partition_sizes, input_size=self.input_size,
            output_size=self.output_size, params_dtype=self.params_dtype,
            weight_loader=self.weight_loader)
        if not reduce_results and (bias and not skip_bias_add):
            raise ValueError("When not reduce the results, adding bias to the results can lead "
                             "to incorrect results")
        if bias:
            self.bias = Parameter(torch.empty(self.output_size, dtype=self.params_dtype))
            set_weight_attrs(self.bias, {"output_dim": 0, "weight_loader": self.weight_loader})
        else:
            self.register_parameter("bias", None)

    def weight_loader(self, param: Parameter, loaded_weight: torch.Tensor):
        input_dim = getattr(param, "input_dim", None)
        if input_dim is not None:
            shard = param.data.shape[input_dim]
            loaded_weight = loaded_weight.narrow(input_dim, self.tp_rank * shard, shard)
        _copy(param.data, loaded_weight)

    def forward(self, input_, input_act: Optional[str] = None, residual: Optional[torch.Tensor] = None,
                want_ssq: bool = False):
        """`input_act="silu_and_mul"`: `input_` is this rank's (gate | up) out of the merged
        column-parallel layer and the layer's input is SiLU(gate) * up (llama.py:89-93) - taken on
        the kernel's x load instead of by an activation launch in between.
        `residual`: the output is all_reduce(y) + residual, i.e. the new residual stream of
        llama.py:276-283; `want_ssq=True` also returns its per-CTA sums of squares for the RMSNorm
        the next column-parallel layer applies on its x load (`input_norm`).  At decode the
        all-reduce, the add and the squares are the table GEMV's epilogue;
        otherwise they run as separate launches."""
        if input_act not in (None, "silu_and_mul"):
            raise ValueError(f"unknown input activation {input_act!r}")
        if self.input_is_parallel:
            input_parallel = input_
        else:
            if input_act is not None:
                raise ValueError("a fused input activation needs input_is_parallel")
            chunk = divide(input_.shape[-1], self.tp_size)
            input_parallel = input_.narrow(-1, self.tp_rank * chunk, chunk).contiguous()
        # bias enters the sum exactly once (linear.py:1206-1208)
        bias_ = None if (self.tp_rank > 0 or self.skip_bias_add) else self.bias
        reduce_ = self.reduce_results and self.tp_size > 1
        ssq = None
        if (reduce_ or residual is not None or want_ssq) and hasattr(self.quant_method, "apply_fused"):
            output, ssq = self.quant_method.apply_fused(self, input_parallel, bias_, residual=residual,
                                                        allreduce=reduce_, silu_mul=input_act is not None,
                                                        want_ssq=want_ssq)
        else:
            if input_act is None:
                output = self.quant_method.apply(self, input_parallel, bias=bias_)
            else:
                output = self.quant_method.apply_silu_mul(self, input_parallel, bias=bias_)
            if reduce_:
                output = tensor_model_parallel_all_reduce(output)
            if residual is not None:
                output = output + residual
        ret = self._ret(output, self.bias if self.skip_bias_add else None)
        if want_ssq:
            return (ret + (ssq,)) if isinstance(ret, tuple) else (ret, ssq)
        return ret
