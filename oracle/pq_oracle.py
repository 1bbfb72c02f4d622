"""CPU oracle for the PQ (AQLM) linear path.  TEST INFRASTRUCTURE ONLY.

This file is a restatement, in numpy/torch on the CPU, of the reference's
torch PQ path (`/root/reference/vllm/model_executor/layers/quantization/aqlm.py`,
cited per function below as ``aqlm.py:<line>``).  It is the checker for the CUDA
kernels in ``vllm_pq_b200/csrc``.  Only ``tests/``, ``__graft_entry__.smoke()``
and ``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs may import it.
The product package ``vllm_pq_b200`` never imports anything from ``oracle/``.

Parity status: PINNED.  ``tests/golden/*.npz`` hold outputs produced by the
reference's own ``dequantize_weight`` / ``generic_dequantize_gemm`` (loaded from
/root/reference by ``oracle/ref_loader.py`` and dumped by
``tests/golden/make_golden.py``); ``tests/test_oracle_golden.py`` checks this
restatement against them bit for bit (weights) and exactly (fp16 layer outputs,
same torch ops).  The reference's own tests hold no numeric vectors for this
path, so the fixtures are the pin.

Conventions: a linear has N outputs, K inputs, P output
partitions; format (b, d, M) = bits per code, subvector length
(``in_group_size``), codebooks per partition.  G = K // d groups per row,
E = 2**b entries per codebook.

    codes      [N, G, M]        int8 (b <= 8) / int16 (b <= 16), signed storage
    codebooks  [P*M, E, 1, d]   fp16 / bf16
    scales     [N, 1, 1, 1]     fp16 / bf16
"""
from __future__ import annotations

from typing import List, Optional, Sequence

import numpy as np
import torch
import torch.nn.functional as F


# --------------------------------------------------------------------------
# storage + unpack                                           aqlm.py:20-34
# --------------------------------------------------------------------------
def get_int_dtype(nbits: int) -> torch.dtype:
    """Storage dtype of the code tensor (aqlm.py:20-29). No sub-byte packing."""
    if nbits <= 8:
        return torch.int8
    if nbits <= 16:
        return torch.int16
    if nbits <= 32:
        return torch.int32
    if nbits <= 64:
        return torch.int64
    raise ValueError(f"No dtype available for {nbits}-bit codebooks")


def unpack_codes_np(codes: np.ndarray, nbits: int) -> np.ndarray:
    """Signed storage -> unsigned index, ``codes.astype(int64) % 2**b``
    (aqlm.py:32-34).  Pure integer arithmetic, so numpy."""
    return np.mod(codes.astype(np.int64), np.int64(1) << nbits)


def unpack_int_data(codes: torch.Tensor, nbits: int) -> torch.Tensor:
    return torch.from_numpy(unpack_codes_np(codes.cpu().numpy(), nbits))


def nbits_of(codebooks: torch.Tensor) -> int:
    """b recovered from the entry count (aqlm.py:86)."""
    return int(codebooks.shape[1]).bit_length() - 1


# --------------------------------------------------------------------------
# dequantisation                                              aqlm.py:37-74
# --------------------------------------------------------------------------
def dequantize_weight(codes_unpacked: torch.Tensor,
                      codebooks: torch.Tensor,
                      scales: Optional[torch.Tensor] = None) -> torch.Tensor:
    """W[n, g*d:(g+1)*d] = scales[n] * sum_m codebooks[m, codes[n, g, m], 0, :].

    Restates aqlm.py:37-74 for out_group_size == 1 (asserted by the reference,
    aqlm.py:178-179).  The reference computes the sum with ``F.embedding_bag(
    mode="sum")`` (aqlm.py:60-65); for fp16 that is "accumulate in fp32 in
    codebook order, round to nearest even once" (checked against the reference
    in tests/test_oracle_golden.py for M up to 8).  It is written out here as
    exactly that, so it also defines the bf16 extension (the CPU
    embedding_bag bf16 kernel does not round to nearest even, and the reference
    does not support bf16 at all - aqlm.py:238-239).
    The scale is then applied in the weight dtype (aqlm.py:70-72): one more
    rounding of the exact product.
    """
    n, g, m = codes_unpacked.shape[-3:]
    m2, e, og, d = codebooks.shape
    assert m == m2 and og == 1
    idx = codes_unpacked.reshape(-1, m).long()
    acc = torch.zeros(idx.shape[0], d, dtype=torch.float32)
    for j in range(m):                       # codebook order, fp32 accumulate
        acc += codebooks[j, :, 0, :].float()[idx[:, j]]
    w = acc.to(codebooks.dtype).view(n, g, 1, d)          # one RNE rounding
    if scales is not None:
        w = w.mul(scales)                                 # aqlm.py:70-72
    return w.swapaxes(-3, -2).reshape(n, g * d)           # aqlm.py:73-74


def dequantize_weight_embedding_bag(codes_unpacked, codebooks, scales=None):
    """The reference's literal formulation (aqlm.py:52-74) kept for the
    cpu_baseline timing leg: same torch ops as the reference, so the timing is
    the reference's."""
    num_out_groups, num_in_groups, num_codebooks = codes_unpacked.shape[-3:]
    num_codebooks, codebook_size, out_group_size, in_group_size = codebooks.shape
    offsets = torch.arange(0, num_codebooks * codebook_size, codebook_size)
    flat = F.embedding_bag(codes_unpacked.flatten(0, -2) + offsets,
                           codebooks.flatten(0, 1).flatten(-2, -1), mode="sum")
    w = flat.view(num_out_groups, num_in_groups, out_group_size, in_group_size)
    if scales is not None:
        w = w.mul(scales)
    return w.swapaxes(-3, -2).reshape(num_out_groups * out_group_size,
                                      num_in_groups * in_group_size)


def _partitions(n_total: int, partition_sizes: Sequence[int]) -> List[int]:
    parts = list(partition_sizes) if len(partition_sizes) else [n_total]
    assert sum(parts) == n_total, (parts, n_total)
    return parts


def dequantize_partitions(codes: torch.Tensor, codebooks: torch.Tensor,
                          partition_sizes: Sequence[int],
                          scales: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Full [N, K] weight of a fused layer: partition p uses rows
    [off_p, off_p + N_p) and codebooks [p*M, (p+1)*M)  (aqlm.py:110-126).
    ``scales=None`` gives what ``aqlm_dequant`` returns (unscaled,
    gemm_kernels.cu:563-592) - parity definition P1; with scales it is P2."""
    nb = nbits_of(codebooks)
    parts = _partitions(codes.shape[0], partition_sizes)
    m = codebooks.shape[0] // len(parts)
    assert m == codes.shape[2]
    out, off = [], 0
    for p, n_p in enumerate(parts):
        c = unpack_int_data(codes[off:off + n_p], nb)
        s = None if scales is None else scales[off:off + n_p]
        out.append(dequantize_weight(c, codebooks[p * m:(p + 1) * m], s))
        off += n_p
    return torch.cat(out, 0)


# --------------------------------------------------------------------------
# layer output                                               aqlm.py:77-127
# --------------------------------------------------------------------------
def generic_dequantize_gemm(x: torch.Tensor, codes: torch.Tensor,
                            codebooks: torch.Tensor, scales: torch.Tensor,
                            partition_sizes: Sequence[int],
                            bias: Optional[torch.Tensor] = None,
                            literal: bool = False) -> torch.Tensor:
    """y = x @ W_scaled^T + bias, per partition, in x's dtype
    (aqlm.py:94-127 calling aqlm.py:77-90).  ``literal=True`` uses the
    embedding_bag formulation (the reference's exact op sequence)."""
    nb = nbits_of(codebooks)
    parts = _partitions(codes.shape[0], partition_sizes)
    m = codebooks.shape[0] // len(parts)
    out = torch.empty(x.shape[:-1] + (scales.shape[0],), dtype=x.dtype)
    deq = dequantize_weight_embedding_bag if literal else dequantize_weight
    off = 0
    for p, n_p in enumerate(parts):
        w = deq(unpack_int_data(codes[off:off + n_p], nb),
                codebooks[p * m:(p + 1) * m], scales[off:off + n_p])
        b = None if bias is None else bias[off:off + n_p]
        out[..., off:off + n_p] = F.linear(x, w, b)
        off += n_p
    return out


def linear_fp32(x: torch.Tensor, codes: torch.Tensor, codebooks: torch.Tensor,
                scales: torch.Tensor, partition_sizes: Sequence[int],
                bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    """The P3 yardstick: the same contraction evaluated
    in fp64 from the *unrounded-scale* operands, i.e. exact sum of codewords
    (as stored), exact scale, exact x - no intermediate rounding at all."""
    nb = nbits_of(codebooks)
    parts = _partitions(codes.shape[0], partition_sizes)
    m = codebooks.shape[0] // len(parts)
    rows, off = [], 0
    for p, n_p in enumerate(parts):
        c = unpack_int_data(codes[off:off + n_p], nb)
        n, g, _ = c.shape
        acc = torch.zeros(n * g, codebooks.shape[3], dtype=torch.float64)
        for j in range(m):
            acc += codebooks[p * m + j, :, 0, :].double()[c.reshape(-1, m)[:, j]]
        w = acc.view(n, g * codebooks.shape[3]) * scales[off:off + n_p].double().view(n, 1)
        rows.append(w)
        off += n_p
    y = x.double() @ torch.cat(rows, 0).t()
    if bias is not None:
        y = y + bias.double()
    return y


def silu_and_mul(x: torch.Tensor) -> torch.Tensor:
    """SiluAndMul.forward_native (vllm/model_executor/layers/activation.py:73-76): the CPU path of
    the MLP's act_fn (llama.py:89-93).  The CUDA kernel (csrc/activation_kernels.cu:35-38) rounds
    the same way - SiLU in fp32 rounded to the dtype, then the product rounded to the dtype - so the
    two agree to the last bit except where expf differs by an ulp across a rounding boundary.
    Pinned bit for bit to the reference's method (tests/golden/layers/silu_*.npz)."""
    d = x.shape[-1] // 2
    return F.silu(x[..., :d]) * x[..., d:]


def fused_add_rms_norm(x: torch.Tensor, residual: Optional[torch.Tensor], weight: torch.Tensor,
                       eps: float):
    """(out, new_residual) of the reference's fused_add_rms_norm / rms_norm kernels
    (csrc/layernorm_kernels.cu:17-44 and :52-98; called from layers/layernorm.py via
    llama.py:272-283):  z = x + residual rounded to the dtype (:75-80, packed half add);
    variance = mean of float(z)^2 (:79,:82-88); out = T(T(float(z) * rstd) * weight) (:91-97: two
    roundings).  The sum of squares is taken in fp64 here - the kernel's fp32 block reduction has no
    fixed order, so rstd may differ from it in the last fp32 bit.
    Pinned (tests/test_oracle_golden.py, tests/golden/layers/norm_*.npz): the new residual is bit-equal
    to RMSNorm.forward_native's (layers/layernorm.py:45-48) and the output is within the two ulps that
    last bit allows of forward_native applied to that rounded residual."""
    z = x if residual is None else (x.float() + residual.float()).to(x.dtype)
    var = z.double().pow(2).mean(-1, keepdim=True)
    rstd = torch.rsqrt(var + eps).float()
    out = ((z.float() * rstd).to(x.dtype).float() * weight.float()).to(x.dtype)
    return out, z


def error_report(y: torch.Tensor, y_ref64: torch.Tensor) -> dict:
    """max-abs, max-abs relative to the largest reference output, rms-relative."""
    diff = (y.double() - y_ref64).abs()
    return {
        "max_abs": float(diff.max()),
        "rel_inf": float(diff.max() / y_ref64.abs().max().clamp_min(1e-30)),
        "rms_rel": float(diff.pow(2).mean().sqrt()
                         / y_ref64.pow(2).mean().sqrt().clamp_min(1e-30)),
    }


# --------------------------------------------------------------------------
# synthetic inputs                           SURVEY.md 8d / benchmark_aqlm.py:266-279
# --------------------------------------------------------------------------
def make_case(n_parts: Sequence[int], k: int, nbits: int, d: int, m: int,
              t: int = 1, dtype: torch.dtype = torch.float16, seed: int = 0,
              bias: bool = False, positive_scales: bool = True) -> dict:
    """Seeded synthetic tensors of the checkpoint layout (aqlm.py:253-316)."""
    g = torch.Generator().manual_seed(seed)
    n = int(sum(n_parts))
    e = 1 << nbits
    lo, hi = -(e // 2), e // 2
    codes = torch.randint(lo, hi, (n, k // d, m), generator=g,
                          dtype=torch.int64).to(get_int_dtype(nbits))
    codebooks = torch.randn(len(n_parts) * m, e, 1, d, generator=g).to(dtype)
    if positive_scales:
        scales = (torch.rand(n, 1, 1, 1, generator=g) + 0.5).to(dtype)
    else:
        scales = torch.randn(n, 1, 1, 1, generator=g).to(dtype)
    x = torch.randn(t, k, generator=g).to(dtype)
    b = torch.randn(n, generator=g).to(dtype) if bias else None
    return dict(codes=codes, codebooks=codebooks, scales=scales, x=x, bias=b,
                partition_sizes=list(n_parts), nbits=nbits, d=d, m=m)
