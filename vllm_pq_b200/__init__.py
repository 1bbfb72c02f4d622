"""vllm_pq_b200 - B200 (sm_100a) product-quantization linear path.

Public surface (mirrors the reference's quantization plugin API for this path):
    PQConfig, PQLinearMethod                     <- AQLMConfig, AQLMLinearMethod
    ops.pq_gemm, ops.pq_dequant                  <- ops.aqlm_gemm, ops.aqlm_dequant
    ColumnParallelLinear, MergedColumnParallelLinear, QKVParallelLinear,
    RowParallelLinear, ReplicatedLinear          <- vllm.model_executor.layers.linear
    tensor_model_parallel_all_reduce             <- vllm.distributed.communication_op
Importing the package loads libpq_b200.so; it fails loudly if the library is
missing (there is no CPU or eager fallback).
"""
from . import _lib

_lib.load()

from . import _custom_ops as ops  # noqa: E402
from .distributed import (destroy_tensor_parallel, get_tensor_model_parallel_rank,  # noqa: E402
                          get_tensor_model_parallel_world_size, init_tensor_parallel,
                          tensor_model_parallel_all_reduce)
from .linear import (ColumnParallelLinear, LinearBase, MergedColumnParallelLinear,  # noqa: E402
                     QKVParallelLinear, ReplicatedLinear, RowParallelLinear)
from .embedding import ParallelLMHead, VocabParallelEmbedding  # noqa: E402
from .moe import PQMixtralMLP, PQMixtralMoE  # noqa: E402
from .plugin_api import (LinearMethodBase, QuantizationConfig, get_quantization_config,  # noqa: E402
                         register_quantization_config)
from .pq import PQConfig, PQLinearMethod, get_int_dtype  # noqa: E402

__all__ = [
    "ops", "PQConfig", "PQLinearMethod", "get_int_dtype", "LinearBase", "ReplicatedLinear",
    "ColumnParallelLinear", "MergedColumnParallelLinear", "QKVParallelLinear", "RowParallelLinear",
    "VocabParallelEmbedding", "ParallelLMHead", "PQMixtralMLP", "PQMixtralMoE", "QuantizationConfig", "LinearMethodBase", "register_quantization_config",
    "get_quantization_config", "init_tensor_parallel", "destroy_tensor_parallel",
    "get_tensor_model_parallel_rank", "get_tensor_model_parallel_world_size",
    "tensor_model_parallel_all_reduce",
]
