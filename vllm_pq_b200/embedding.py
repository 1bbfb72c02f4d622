"""Vocabulary layers over PQ weights: VocabParallelEmbedding and ParallelLMHead.

Condensed mirror of vllm/model_executor/layers/vocab_parallel_embedding.py (the reference keeps both
dense: aqlm.py:213-217; a PQ config with ``lm_head_quantized`` hands them the PQ method instead):
  * the vocabulary is padded to a multiple of 64 and split by rank along dim 0 (:60-66, :229-246);
  * ``create_weights`` is called like for a column-parallel linear - input = embedding_dim, one
    output partition of ``num_embeddings_per_partition`` rows (:263-270) - so the checkpoint tensors
    are the PQ triple ``codes [V/tp, H/d, M]``, ``codebooks [M, E, 1, d]``, ``scales [V/tp, 1, 1, 1]``;
  * forward: ids of other shards are masked, looked up through ``quant_method.embedding`` (the PQ
    row-gather kernel), zeroed, and all-reduced over the TP group (:395-411);
  * ParallelLMHead is the same parameters used through ``quant_method.apply`` (logits = h . W^T,
    :414-470 and logits_processor.py:_get_logits): one more PQ GEMV / GEMM with N = vocabulary.
Added-vocabulary (LoRA) bookkeeping is out of scope.
"""
from __future__ import annotations

from typing import Optional

import torch
from torch import nn
from torch.nn import Parameter

from .distributed import (get_tensor_model_parallel_rank, get_tensor_model_parallel_world_size,
                          tensor_model_parallel_all_reduce)
from .plugin_api import QuantizationConfig, set_weight_attrs

DEFAULT_VOCAB_PADDING_SIZE = 64


def pad_vocab_size(vocab_size: int, pad_to: int = DEFAULT_VOCAB_PADDING_SIZE) -> int:
    return ((vocab_size + pad_to - 1) // pad_to) * pad_to


class VocabParallelEmbedding(nn.Module):
    def __init__(self, num_embeddings: int, embedding_dim: int, params_dtype: Optional[torch.dtype] = None,
                 padding_size: int = DEFAULT_VOCAB_PADDING_SIZE, quant_config: Optional[QuantizationConfig] = None,
                 prefix: str = ""):
        super().__init__()
        self.tp_rank = get_tensor_model_parallel_rank()
        self.tp_size = get_tensor_model_parallel_world_size()
        self.num_embeddings = num_embeddings
        self.embedding_dim = embedding_dim
        self.num_embeddings_padded = pad_vocab_size(num_embeddings, padding_size)
        if self.num_embeddings_padded % self.tp_size:
            raise ValueError("padded vocabulary is not divisible by the tensor-parallel size")
        self.num_embeddings_per_partition = self.num_embeddings_padded // self.tp_size
        self.vocab_start = self.tp_rank * self.num_embeddings_per_partition
        self.vocab_end = self.vocab_start + self.num_embeddings_per_partition
        if quant_config is None:
            raise ValueError("vllm_pq_b200 layers need a quantization config (PQConfig)")
        method = quant_config.get_quant_method(self, prefix=prefix)
        if method is None:
            raise NotImplementedError("this PQConfig keeps the vocabulary layers dense "
                                      "(lm_head_quantized is false); the dense layers are vLLM's own")
        if not callable(getattr(method, "embedding", None)):
            raise NotImplementedError(f"The class {type(method).__name__} must implement the 'embedding' method")
        self.quant_method = method
        self.params_dtype = params_dtype if params_dtype is not None else torch.get_default_dtype()
        self.quant_method.create_weights(self, self.embedding_dim, [self.num_embeddings_per_partition],
                                         self.embedding_dim, self.num_embeddings_padded,
                                         params_dtype=self.params_dtype, weight_loader=self.weight_loader)

    def weight_loader(self, param: Parameter, loaded_weight: torch.Tensor):
        """Rows [vocab_start, vocab_end) of a full-vocabulary tensor; rows past the checkpoint's
        vocabulary (padding) are zero - a zero scale makes a padded row dequantise to zeros
        (vocab_parallel_embedding.py:361-393)."""
        output_dim = getattr(param, "output_dim", None)
        if output_dim is None:                                 # codebooks: replicated
            assert param.data.shape == loaded_weight.shape
            param.data.copy_(loaded_weight)
            return
        have = max(0, min(loaded_weight.shape[output_dim], self.vocab_end) - self.vocab_start)
        param.data.zero_()
        if have:
            param.data.narrow(output_dim, 0, have).copy_(loaded_weight.narrow(output_dim, self.vocab_start, have))

    def forward(self, input_: torch.Tensor) -> torch.Tensor:
        if self.tp_size > 1:
            mask = (input_ < self.vocab_start) | (input_ >= self.vocab_end)
            local = (input_ - self.vocab_start).masked_fill(mask, -1)     # -1: the kernel writes a zero row
        else:
            local = input_
        out = self.quant_method.embedding(self, local.long())
        if self.tp_size > 1:
            out = tensor_model_parallel_all_reduce(out)
        return out


class ParallelLMHead(VocabParallelEmbedding):
    """Output embedding: same parameters, used as a linear layer by the logits processor
    (vocab_parallel_embedding.py:414-470; logits_processor.py: ``lm_head.quant_method.apply``)."""

    def __init__(self, num_embeddings: int, embedding_dim: int, bias: bool = False,
                 params_dtype: Optional[torch.dtype] = None, padding_size: int = DEFAULT_VOCAB_PADDING_SIZE,
                 quant_config: Optional[QuantizationConfig] = None, prefix: str = ""):
        super().__init__(num_embeddings, embedding_dim, params_dtype, padding_size, quant_config, prefix)
        if bias:
            self.bias = Parameter(torch.empty(self.num_embeddings_per_partition, dtype=self.params_dtype))
            set_weight_attrs(self.bias, {"output_dim": 0, "weight_loader": self.weight_loader})
        else:
            self.register_parameter("bias", None)

    def tie_weights(self, embed_tokens: VocabParallelEmbedding):
        for name in ("codes", "codebooks", "scales"):
            setattr(self, name, getattr(embed_tokens, name))
        return self

    def compute_logits(self, hidden_states: torch.Tensor) -> torch.Tensor:
        """This rank's slice of the logits, [T, V_padded / tp] (the gather over ranks and the removal of
        the padding belong to the logits processor, out of scope)."""
        return self.quant_method.apply(self, hidden_states, self.bias)

    def forward(self, input_):
        raise RuntimeError("LMHead's weights should be used in the sampler.")     # as the reference
