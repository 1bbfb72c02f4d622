"""Per-expert PQ linears: the quantised Mixtral MoE block.

Mirror of the reference's NON-fused quantised Mixtral (vllm/model_executor/models/mixtral_quant.py:58-160),
the path vLLM routes an AQLM Mixtral checkpoint to (model_loader/utils.py:89-99): every expert is three
``ReplicatedLinear`` layers (w1, w3: hidden -> ffn; w2: ffn -> hidden) that take the PQ method through the
ordinary plugin boundary, experts are split between the tensor-parallel ranks, every local expert runs on
ALL tokens and is weighted by its (possibly zero) routing weight, and the ranks' sums are all-reduced.
Nothing here is a new kernel: an expert is ``pq::gemm`` three times.  Differences from the reference, all
host-side: the router stays dense (as there: ``quant_config=None``) but is a plain parameter, since this
package has no unquantised linear method; an expert none of the tokens routed to is skipped at decode
(its contribution is exactly zero) instead of being multiplied by zero.
"""
from __future__ import annotations

from typing import Optional

import numpy as np
import torch
import torch.nn.functional as F
from torch import nn

from .distributed import (get_tensor_model_parallel_rank, get_tensor_model_parallel_world_size,
                          tensor_model_parallel_all_reduce)
from .linear import ReplicatedLinear
from .plugin_api import QuantizationConfig


class PQMixtralMLP(nn.Module):
    """One expert: w2(SiLU(w1 x) * w3 x)  (mixtral_quant.py:58-95)."""

    def __init__(self, hidden_size: int, intermediate_size: int, params_dtype: Optional[torch.dtype] = None,
                 quant_config: Optional[QuantizationConfig] = None):
        super().__init__()
        kw = dict(bias=False, params_dtype=params_dtype, quant_config=quant_config)
        self.w1 = ReplicatedLinear(hidden_size, intermediate_size, **kw)
        self.w2 = ReplicatedLinear(intermediate_size, hidden_size, **kw)
        self.w3 = ReplicatedLinear(hidden_size, intermediate_size, **kw)

    def forward(self, hidden_states: torch.Tensor) -> torch.Tensor:
        w1_out, _ = self.w1(hidden_states)
        w3_out, _ = self.w3(hidden_states)
        out, _ = self.w2(F.silu(w1_out) * w3_out)
        return out


class PQMixtralMoE(nn.Module):
    """Router + this rank's experts + all-reduce  (mixtral_quant.py:98-160)."""

    def __init__(self, num_experts: int, top_k: int, hidden_size: int, intermediate_size: int,
                 params_dtype: Optional[torch.dtype] = None, quant_config: Optional[QuantizationConfig] = None):
        super().__init__()
        self.rank = get_tensor_model_parallel_rank()
        self.tp_size = get_tensor_model_parallel_world_size()
        self.num_total_experts, self.top_k = num_experts, top_k
        if self.tp_size > num_experts:
            raise ValueError(f"Tensor parallel size {self.tp_size} is greater than the number of experts {num_experts}.")
        self.expert_indicies = np.array_split(range(num_experts), self.tp_size)[self.rank].tolist()
        if not self.expert_indicies:
            raise ValueError(f"Rank {self.rank} has no experts assigned to it.")
        self.experts = nn.ModuleList([
            PQMixtralMLP(hidden_size, intermediate_size, params_dtype, quant_config) if idx in self.expert_indicies else None
            for idx in range(num_experts)])
        dtype = params_dtype if params_dtype is not None else torch.get_default_dtype()
        self.gate_weight = nn.Parameter(torch.empty(num_experts, hidden_size, dtype=dtype), requires_grad=False)

    def forward(self, hidden_states: torch.Tensor) -> torch.Tensor:
        num_tokens, hidden_dim = hidden_states.shape
        router_logits = F.linear(hidden_states, self.gate_weight)
        routing_weights = F.softmax(router_logits, dim=1, dtype=torch.float)
        routing_weights, selected_experts = torch.topk(routing_weights, self.top_k, dim=-1)
        routing_weights /= routing_weights.sum(dim=-1, keepdim=True)
        final = None
        # which local experts got any token: one small device-to-host read per step at decode buys
        # skipping the (num_local - top_k) experts whose weight is exactly zero for every token
        used = None
        if num_tokens <= 8 and not (hidden_states.is_cuda and torch.cuda.is_current_stream_capturing()):
            used = set(selected_experts.flatten().tolist())
        for expert_idx in self.expert_indicies:
            if used is not None and expert_idx not in used:
                continue
            expert_mask = (selected_experts == expert_idx)
            expert_weights = (routing_weights * expert_mask).sum(dim=-1, keepdim=True)
            current = self.experts[expert_idx](hidden_states).mul_(expert_weights)
            final = current if final is None else final.add_(current)
        if final is None:
            final = torch.zeros_like(hidden_states)
        return tensor_model_parallel_all_reduce(final).view(num_tokens, hidden_dim)
