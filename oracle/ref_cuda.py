"""The reference's own CUDA PQ (AQLM) kernels on the GPU box.  TEST INFRASTRUCTURE ONLY.

``oracle/Makefile`` compiles /root/reference/csrc/quantization/aqlm/gemm_kernels.cu (unmodified,
from where it lies) for sm_100a into ``oracle/_ref/libref_aqlm.so`` and registers the reference's
two ops under the namespace ``pqref`` (schemas of torch_bindings.cpp:161-173).  Used by
``tests/test_gpu_ref_cuda.py`` (parity gate P4) and ``tools/microbench.py --ref``
(the same-box "kernel to beat").  Formats: 1x16 and 2x8 with d = 8, fp16 only
(gemm_kernels.cu:527-540).  Nothing in ``vllm_pq_b200`` imports this.
"""
from __future__ import annotations

import os

import torch

LIB = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "libref_aqlm.so")
_loaded = False


def available() -> bool:
    return os.path.isfile(LIB) and torch.cuda.is_available()


def load() -> None:
    global _loaded
    if not _loaded:
        torch.ops.load_library(LIB)
        _loaded = True


def supports(nbits: int, d: int, m: int, dtype: torch.dtype) -> bool:
    """gemm_kernels.cu:527-540 / 568-592: 1x16 and 2x8, d = 8, half only."""
    return dtype == torch.float16 and d == 8 and ((nbits == 16 and m == 1) or (nbits == 8 and m == 2))


def aqlm_gemm(x, codes, codebooks, scales, parts, bias=None):
    load()
    return torch.ops.pqref.aqlm_gemm(x, codes, codebooks, scales, list(parts), bias)


def aqlm_dequant(codes, codebooks, parts):
    load()
    return torch.ops.pqref.aqlm_dequant(codes, codebooks, list(parts))


def optimized_dequantize_gemm(x, codes, codebooks, scales, parts, bias=None):
    """aqlm.py:132-156 over the reference's CUDA dequant (what the reference runs for > 6 tokens)."""
    import torch.nn.functional as F
    w = aqlm_dequant(codes, codebooks, parts)
    if bias is None:
        y = F.linear(x, w, None)
        y *= scales.flatten().view((1,) * (y.dim() - 1) + (-1,))
        return y
    w = w * scales.view(-1, 1)
    return F.linear(x, w, bias)
