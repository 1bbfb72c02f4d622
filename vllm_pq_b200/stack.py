"""The PQ linears of a Llama decoder stack, wired the way the reference wires them.

Only the four projections of each layer (qkv_proj, o_proj, gate_up_proj,
down_proj - llama.py:70-83,140-156 in the reference) with vLLM tensor parallelism:
column-parallel QKV and gate/up, row-parallel O and down, one all-reduce after each
row-parallel layer.  Attention, norms, rotary and the activation are NOT here (out of
scope, SURVEY.md 8d "Tokens/s"); between two linears the activations are simply
narrowed to the next layer's input width, so every launch still depends on the previous
one exactly as in a model.  Used by bench.py and the GPU tests.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Tuple

import torch
from torch import nn

from .distributed import get_tensor_model_parallel_world_size
from .linear import MergedColumnParallelLinear, QKVParallelLinear, RowParallelLinear
from .pq import PQConfig


@dataclass
class LlamaShape:
    name: str
    hidden: int
    intermediate: int
    heads: int
    kv_heads: int
    layers: int

    @property
    def head_dim(self) -> int:
        return self.hidden // self.heads


LLAMA2_7B = LlamaShape("llama-2-7b", 4096, 11008, 32, 32, 32)
LLAMA3_8B = LlamaShape("llama-3-8b", 4096, 14336, 32, 8, 32)
LLAMA2_70B = LlamaShape("llama-2-70b", 8192, 28672, 64, 8, 80)
SHAPES = {s.name: s for s in (LLAMA2_7B, LLAMA3_8B, LLAMA2_70B)}


class PQDecoderLinears(nn.Module):
    """qkv -> o -> gate_up -> down of one decoder layer."""

    def __init__(self, shape: LlamaShape, cfg: PQConfig, dtype: torch.dtype, fused_act: bool = False):
        super().__init__()
        self.fused_act = fused_act
        kw = dict(bias=False, params_dtype=dtype, quant_config=cfg)
        self.qkv_proj = QKVParallelLinear(shape.hidden, shape.head_dim, shape.heads, shape.kv_heads, **kw)
        self.o_proj = RowParallelLinear(shape.heads * shape.head_dim, shape.hidden, **kw)
        self.gate_up_proj = MergedColumnParallelLinear(shape.hidden, [shape.intermediate] * 2, **kw)
        self.down_proj = RowParallelLinear(shape.intermediate, shape.hidden, **kw)
        self.q_width = self.qkv_proj.num_heads * shape.head_dim        # per-rank attention width
        self.mlp_width = shape.intermediate // get_tensor_model_parallel_world_size()

    def linears(self) -> List[Tuple[str, nn.Module]]:
        return [("qkv_proj", self.qkv_proj), ("o_proj", self.o_proj),
                ("gate_up_proj", self.gate_up_proj), ("down_proj", self.down_proj)]

    def forward(self, h: torch.Tensor) -> torch.Tensor:
        qkv, _ = self.qkv_proj(h)
        attn = qkv[..., :self.q_width]                  # stand-in for attention: keeps the dependency
        h, _ = self.o_proj(attn)
        gate_up, _ = self.gate_up_proj(h)
        if self.fused_act:                              # the real MLP: SiLU(gate) * up on down_proj's x load
            h, _ = self.down_proj(gate_up, input_act="silu_and_mul")
            return h
        act = gate_up[..., :self.mlp_width]             # stand-in for SiLU(gate) * up
        h, _ = self.down_proj(act)
        return h


class PQLlamaLinearStack(nn.Module):
    def __init__(self, shape: LlamaShape, cfg: PQConfig, dtype: torch.dtype = torch.float16,
                 n_layers: int = None, fused_act: bool = False):
        super().__init__()
        self.shape, self.cfg, self.dtype = shape, cfg, dtype
        n_layers = shape.layers if n_layers is None else n_layers
        self.layers = nn.ModuleList(PQDecoderLinears(shape, cfg, dtype, fused_act) for _ in range(n_layers))

    @torch.no_grad()
    def init_synthetic(self, seed: int = 0) -> "PQLlamaLinearStack":
        """Uniform random codes over all entries, signed storage (benchmark_aqlm.py:266-273);
        Gaussian codebooks, zero-mean over the entries and scaled so activations stay O(1) through
        the stack; scales uniform with E[s^2] = 1.

        Zero mean matters: every row of a PQ weight draws from the same E codewords, so the mean
        codeword adds the same <cbar, sum_g x_g> to every output.  That common mode feeds the next
        layer's sum_g x_g and is multiplied by G * |cbar . 1| ~ 1.4 (K = 4096) or 2.3 (K = 11008) per
        linear: a random walk that overflows fp16 within a few dozen layers (the r1 stack ended
        non-finite for exactly this reason, profiles/r1_34_bench_act.json "slice_finite")."""
        dev = next(self.parameters()).device
        g = torch.Generator(device=dev).manual_seed(seed)
        e = 1 << self.cfg.nbits_per_codebook
        s_norm = (1.0 + 1.0 / 12.0) ** -0.5                # rand + 0.5 has E[s^2] = 13/12
        for layer in self.layers:
            for _, lin in layer.linears():
                k = lin.codes.shape[1] * self.cfg.in_group_size * (
                    get_tensor_model_parallel_world_size() if isinstance(lin, RowParallelLinear) else 1)
                lin.codes.data.copy_(torch.randint(-(e // 2), e // 2, lin.codes.shape, device=dev,
                                                   generator=g, dtype=torch.int32))
                std = (1.0 / (k * self.cfg.num_codebooks)) ** 0.5
                cb = torch.randn(lin.codebooks.shape, device=dev, generator=g) * std
                lin.codebooks.data.copy_(cb - cb.mean(dim=1, keepdim=True))
                lin.scales.data.copy_((torch.rand(lin.scales.shape, device=dev, generator=g) + 0.5) * s_norm)
                lin.quant_method.process_weights_after_loading(lin)
        return self

    def forward(self, h: torch.Tensor) -> torch.Tensor:
        for layer in self.layers:
            h = layer(h)
        return h

    # ---- bookkeeping for the roofline -----------------------------------------------------
    def code_bytes(self) -> int:
        return sum(p.numel() * p.element_size() for n, p in self.named_parameters() if n.endswith("codes"))

    def launches_per_step(self) -> int:
        tp = get_tensor_model_parallel_world_size()
        per_layer = 4 + (2 if tp > 1 else 0)
        return per_layer * len(self.layers)


def algorithmic_bytes(lin: nn.Module, t: int) -> int:
    """every tensor of one call counted once, workspaces excluded."""
    n, g, m = lin.codes.shape
    d = lin.codebooks.shape[3]
    return (lin.codes.numel() * lin.codes.element_size() + lin.codebooks.numel() * 2 + n * 2
            + t * g * d * 2 + t * n * 2)
