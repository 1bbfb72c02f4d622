"""torch custom ops ``pq::gemm`` / ``pq::dequant`` over the C ABI.

Mirror of the reference's op boundary for this path:
  torch.ops._C.aqlm_gemm / aqlm_dequant   csrc/torch_bindings.cpp:161-173
  python wrappers                          vllm/_custom_ops.py:646-658
  fake (meta) implementations              vllm/_custom_ops.py:347-371
Same argument order and meaning, same output shape rules, RuntimeError for an
unsupported request (TORCH_CHECK in gemm_kernels.cu:539-540).  Differences:
``dequant`` derives in_features from the codebook's subvector length instead of
hard-coding 8 (gemm_kernels.cu:557), and both accept bf16.

The CUDA implementation only reads its inputs and returns a fresh tensor
(SURVEY.md 8b "Ownership"); it launches on torch's current stream, does not
synchronise and does not allocate outside torch's allocator, so it can be
captured in a CUDA graph (model_runner.py:1428-1562 in the reference).
"""
from __future__ import annotations

import ctypes
from typing import List, Optional

import torch

from . import _lib

_DTYPES = {torch.float16: _lib.PQ_F16, torch.bfloat16: _lib.PQ_BF16}

_LIBRARY = torch.library.Library("pq", "DEF")
_LIBRARY.define("gemm(Tensor input, Tensor codes, Tensor codebooks, Tensor scales, "
                "int[] codebook_partition_sizes, Tensor? bias) -> Tensor")
_LIBRARY.define("gemm_algo(Tensor input, Tensor codes, Tensor codebooks, Tensor scales, "
                "int[] codebook_partition_sizes, Tensor? bias, int algo) -> Tensor")
_LIBRARY.define("gemm_silu_mul(Tensor input, Tensor codes, Tensor codebooks, Tensor scales, "
                "int[] codebook_partition_sizes, Tensor? bias, int algo) -> Tensor")
_LIBRARY.define("silu_and_mul(Tensor input) -> Tensor")
_LIBRARY.define("gemm_fused(Tensor input, Tensor codes, Tensor codebooks, Tensor scales, "
                "int[] codebook_partition_sizes, Tensor? bias, Tensor? residual, Tensor? norm_ssq, "
                "Tensor? norm_weight, float norm_eps, bool allreduce, bool silu_mul, bool want_ssq, "
                "Tensor? codes_padded) -> (Tensor, Tensor)")
_LIBRARY.define("fused_add_rms_norm(Tensor(a!) input, Tensor(b!)? residual, Tensor weight, float epsilon) -> ()")
_LIBRARY.define("embedding(Tensor ids, Tensor codes, Tensor codebooks, Tensor scales) -> Tensor")
_LIBRARY.define("dequant(Tensor codes, Tensor codebooks, int[] codebook_partition_sizes) -> Tensor")
_LIBRARY.define("dequant_scaled(Tensor codes, Tensor codebooks, Tensor scales, "
                "int[] codebook_partition_sizes) -> Tensor")


# One zero-initialised workspace per (device, stream).  The split-K kernels keep ticket counters in
# its first 64 KiB and leave them zero (include/pq_b200.h), so it is zeroed only when allocated.
#  * per stream: two streams of one device must not share tickets and partial sums;
#  * never freed: a CUDA graph captured earlier holds the raw pointer, so a buffer that is outgrown
#    is retired (kept alive), not dropped - replaying the old graph stays valid;
#  * the SiLU*mul scratch of gemm_silu_mul is NOT in here: it is a per-call torch tensor
#    (stream-ordered, graph-pool safe), see _gemm_impl.
_WORKSPACES = {}
_RETIRED = []
_WS_MIN_BYTES = 8 << 20


def _workspace(device: torch.device, nbytes: int) -> Optional[torch.Tensor]:
    key = (device.index if device.index is not None else torch.cuda.current_device(),
           torch.cuda.current_stream(device).cuda_stream)
    ws = _WORKSPACES.get(key)
    if ws is not None and ws.numel() >= nbytes:
        return ws
    if torch.cuda.is_current_stream_capturing():
        # No buffer of this stream from before the capture (callers warm up eagerly on the capture
        # stream first, as vLLM does): allocating here would put a memset node in front of every
        # launch.  The workspace is optional (include/pq_b200.h): without it the table GEMV plans
        # without K groups.
        return None
    size = max(int(nbytes) * 2, _WS_MIN_BYTES)
    new = torch.zeros(size, dtype=torch.uint8, device=device)
    if ws is not None:
        _RETIRED.append(ws)
    _WORKSPACES[key] = new
    return new


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def _format(codes: torch.Tensor, codebooks: torch.Tensor, parts: List[int]):
    if codes.dim() != 3 or codebooks.dim() != 4:
        raise RuntimeError("PQ: codes must be [N, G, M] and codebooks [P*M, E, 1, d]")
    if codebooks.dtype not in _DTYPES:
        raise RuntimeError(f"PQ: codebooks dtype {codebooks.dtype} not supported (fp16/bf16)")
    parts = list(parts) if len(parts) else [codes.shape[0]]
    n, g, m = codes.shape
    pm, e, og, d = codebooks.shape
    if og != 1:
        raise RuntimeError("PQ: out_group_size must be 1")          # aqlm.py:178-179
    if pm != m * len(parts) or (e & (e - 1)) != 0:
        raise RuntimeError(f"PQ with {pm / len(parts)} codebooks and {e} entries is not "
                           "currently supported.")                 # gemm_kernels.cu:539
    nbits = e.bit_length() - 1
    want = torch.int8 if nbits <= 8 else torch.int16
    if codes.dtype != want:
        raise RuntimeError(f"PQ: {nbits}-bit codes must be stored as {want}, got {codes.dtype}")
    if not codes.is_contiguous() or not codebooks.is_contiguous():
        raise RuntimeError("PQ: codes and codebooks must be contiguous")
    return parts, n, g, m, d, nbits


def _require_cuda(t: torch.Tensor):
    if not t.is_cuda:
        raise RuntimeError("PQ kernels are CUDA (sm_100a) only; there is no CPU path")


def _dequant_impl(codes, codebooks, scales, parts):
    _require_cuda(codes)
    lib = _lib.load()
    parts, n, g, m, d, nbits = _format(codes, codebooks, parts)
    if scales is not None:
        if scales.dtype != codebooks.dtype or scales.numel() != n or not scales.is_contiguous():
            raise RuntimeError("PQ: scales must be contiguous [N,1,1,1] of the codebook dtype")
    out = torch.empty((n, g * d), dtype=codebooks.dtype, device=codebooks.device)
    arr, np_ = _lib.parts_array(parts)
    with torch.cuda.device(codes.device):
        rc = lib.pq_b200_dequant(codes.data_ptr(), codebooks.data_ptr(), _ptr(scales), out.data_ptr(),
                                 n, g * d, d, nbits, m, arr, np_, _DTYPES[codebooks.dtype],
                                 torch.cuda.current_stream().cuda_stream)
    _lib.check(rc, "pq::dequant")
    return out


def _gemm_impl(input, codes, codebooks, scales, parts, bias, algo, silu_mul=False):
    _require_cuda(input)
    lib = _lib.load()
    parts, n, g, m, d, nbits = _format(codes, codebooks, parts)
    k = g * d
    kx = 2 * k if silu_mul else k              # silu_mul: input is (gate | up), llama.py:89-93
    if input.shape[-1] != kx:
        raise RuntimeError(f"PQ: input has {input.shape[-1]} features, weight has {k}"
                           + (" (expects gate|up, twice that)" if silu_mul else ""))
    dt = input.dtype
    if dt not in _DTYPES or codebooks.dtype != dt or scales.dtype != dt:
        raise RuntimeError("PQ: input, codebooks and scales must share one of fp16/bf16")
    if scales.numel() != n or not scales.is_contiguous():
        raise RuntimeError("PQ: scales must be contiguous [N,1,1,1]")
    if bias is not None and (bias.dtype != dt or bias.numel() != n or not bias.is_contiguous()):
        raise RuntimeError("PQ: bias must be contiguous [N] of the input dtype")
    x = input.reshape(-1, kx)
    if x.stride(-1) != 1 or (x.stride(0) * 2) % 16 or x.data_ptr() % 16:
        x = x.contiguous()
    t = x.shape[0]
    out = torch.empty((t, n), dtype=dt, device=input.device)
    if t == 0:
        return out.reshape(input.shape[:-1] + (n,))
    arr, np_ = _lib.parts_array(parts)
    with torch.cuda.device(input.device):
        if algo == _lib.ALGO_AUTO:
            algo = lib.pq_b200_pick_algo(t, n, k, d, nbits, m)
        if silu_mul and not (algo == _lib.ALGO_GEMV_LUT_TMA and
                             lib.pq_b200_algo_supported(algo, t, n, k, d, nbits, m)):
            # no kernel of this family takes the activation on its x load: one silu_and_mul launch
            # into a per-call tensor, then the plain linear (bit-identical to the C entry point's
            # own two-launch form, tests/test_gpu_act.py)
            act = _silu_and_mul_cuda(x)
            return _gemm_impl(act, codes, codebooks, scales, parts, bias, algo).reshape(input.shape[:-1] + (n,))
        ws_bytes = lib.pq_b200_workspace_bytes(algo, t, n, k, d, nbits, m, np_)
        ws = _workspace(input.device, ws_bytes) if ws_bytes else None
        fn = lib.pq_b200_gemm_silu_mul if silu_mul else lib.pq_b200_gemm
        rc = fn(x.data_ptr(), x.stride(0), codes.data_ptr(), codebooks.data_ptr(),
                scales.data_ptr(), _ptr(bias), out.data_ptr(), n, t, n, k, d, nbits, m,
                arr, np_, _DTYPES[dt], algo, _ptr(ws), ws.numel() if ws is not None else 0,
                torch.cuda.current_stream().cuda_stream)
    _lib.check(rc, "pq::gemm")
    return out.reshape(input.shape[:-1] + (n,))


def gemm_fused_supported(t: int, n: int, k: int, d: int, nbits: int, m: int) -> bool:
    """Does the decode table GEMV - the kernel that carries the fused epilogue - cover this problem?"""
    return bool(_lib.load().pq_b200_gemm_fused_supported(t, n, k, d, nbits, m))


def _gemm_fused_cuda(input, codes, codebooks, scales, parts, bias, residual, norm_ssq, norm_weight, norm_eps,
                     allreduce, silu_mul, want_ssq, codes_padded=None):
    """pq_b200_gemm_fused: the table GEMV with (norm on the x load | SiLU*mul on the x load), and in its
    epilogue (all-reduce over the TP group,) (residual add,) (sum of squares per CTA).  SURVEY 8f N1."""
    _require_cuda(input)
    lib = _lib.load()
    parts, n, g, m, d, nbits = _format(codes, codebooks, parts)
    k = g * d
    kx = 2 * k if silu_mul else k
    dt = input.dtype
    if input.shape[-1] != kx or dt not in _DTYPES or codebooks.dtype != dt or scales.dtype != dt:
        raise RuntimeError("PQ: gemm_fused input/weight mismatch")
    x = input.reshape(-1, kx)
    if x.stride(-1) != 1 or (x.stride(0) * 2) % 16 or x.data_ptr() % 16:
        x = x.contiguous()
    t = x.shape[0]
    if t < 1 or not gemm_fused_supported(t, n, k, d, nbits, m):
        raise RuntimeError("PQ: gemm_fused needs the decode table GEMV (1..8 tokens, 8-bit codes); "
                           "run the steps separately")
    ep = _lib.Epilogue()
    keep = []
    if allreduce:
        from . import distributed
        comm = distributed.fused_allreduce_comm()
        if comm is None or t * n > comm.LL_SLOT_ELEMS:
            raise RuntimeError("PQ: gemm_fused(allreduce=True) without a peer-memory communicator")
        comm.fill_epilogue(ep)
    if residual is not None:
        r2 = residual.reshape(-1, n)
        if r2.dtype != dt or r2.shape[0] != t or r2.stride(-1) != 1:
            raise RuntimeError("PQ: residual must be [T, N] of the input dtype")
        ep.residual, ep.residual_stride = r2.data_ptr(), r2.stride(0)
        keep.append(r2)
    if norm_weight is not None:
        if silu_mul or norm_ssq is None or norm_weight.dtype != dt or norm_weight.numel() != k:
            raise RuntimeError("PQ: norm on load needs norm_ssq [T, slots] fp32 and a [K] weight")
        nss = norm_ssq.reshape(t, -1)
        if nss.dtype != torch.float32 or not nss.is_contiguous() or not norm_weight.is_contiguous():
            raise RuntimeError("PQ: norm_ssq must be contiguous fp32")
        ep.norm_ssq, ep.norm_weight = nss.data_ptr(), norm_weight.data_ptr()
        ep.norm_slots, ep.norm_eps = nss.shape[1], float(norm_eps)
        keep.append(nss)
    ep.input_silu_mul = 1 if silu_mul else 0
    if codes_padded is not None:
        if (codes_padded.dim() != 2 or codes_padded.shape[0] != n or codes_padded.shape[1] < g * m
                or codes_padded.element_size() != 1 or not codes_padded.is_contiguous()):
            raise RuntimeError("PQ: codes_padded must be a contiguous [N, pitch] byte tensor, pitch >= G * M")
        ep.codes_padded, ep.codes_pitch = codes_padded.data_ptr(), codes_padded.shape[1]
    arr, np_ = _lib.parts_array(parts)
    out = torch.empty((t, n), dtype=dt, device=input.device)
    with torch.cuda.device(input.device):
        ws_bytes = lib.pq_b200_workspace_bytes(_lib.ALGO_GEMV_LUT_TMA, t, n, k, d, nbits, m, np_)
        ws = _workspace(input.device, ws_bytes) if ws_bytes else None
        ws_n = ws.numel() if ws is not None else 0
        if want_ssq:
            slots = lib.pq_b200_gemm_fused_slots(t, n, k, d, nbits, m, arr, np_, ws_n)
            _lib.check(0 if slots > 0 else int(slots), "pq::gemm_fused (slots)")
            ssq = torch.empty((t, slots), dtype=torch.float32, device=input.device)
            ep.ssq_out, ep.ssq_slots = ssq.data_ptr(), slots
        else:
            ssq = torch.empty((0,), dtype=torch.float32, device=input.device)
        rc = lib.pq_b200_gemm_fused(x.data_ptr(), x.stride(0), codes.data_ptr(), codebooks.data_ptr(),
                                    scales.data_ptr(), _ptr(bias), out.data_ptr(), n, t, n, k, d, nbits, m,
                                    arr, np_, _DTYPES[dt], _ptr(ws), ws_n,
                                    torch.cuda.current_stream().cuda_stream, ctypes.byref(ep))
    _lib.check(rc, "pq::gemm_fused")
    del keep
    return out.reshape(input.shape[:-1] + (n,)), ssq


def _fused_add_rms_norm_cuda(input, residual, weight, epsilon):
    _require_cuda(input)
    h = input.shape[-1]
    x = input.view(-1, h)
    if input.dtype not in _DTYPES or weight.dtype != input.dtype or weight.numel() != h or x.stride(-1) != 1:
        raise RuntimeError("PQ: fused_add_rms_norm takes fp16/bf16 [..., H] and a [H] weight")
    r = None
    if residual is not None:
        r = residual.view(-1, h)
        if r.dtype != input.dtype or r.shape != x.shape or r.stride(-1) != 1:
            raise RuntimeError("PQ: residual must match input")
    if x.shape[0]:
        with torch.cuda.device(input.device):
            rc = _lib.load().pq_b200_fused_add_rms_norm(x.data_ptr(), x.stride(0), _ptr(r), r.stride(0) if r is not None else 0,
                                                        weight.data_ptr(), float(epsilon), x.shape[0], h,
                                                        _DTYPES[input.dtype], torch.cuda.current_stream().cuda_stream)
        _lib.check(rc, "pq::fused_add_rms_norm")


def _embedding_cuda(ids, codes, codebooks, scales):
    _require_cuda(codes)
    lib = _lib.load()
    parts, n, g, m, d, nbits = _format(codes, codebooks, [])
    if ids.dtype != torch.int64 or not ids.is_cuda:
        raise RuntimeError("PQ: embedding ids must be an int64 CUDA tensor")
    if scales.dtype != codebooks.dtype or scales.numel() != n or not scales.is_contiguous():
        raise RuntimeError("PQ: scales must be contiguous [N,1,1,1] of the codebook dtype")
    flat = ids.reshape(-1).contiguous()
    out = torch.empty((flat.numel(), g * d), dtype=codebooks.dtype, device=codebooks.device)
    with torch.cuda.device(codes.device):
        rc = lib.pq_b200_embedding(flat.data_ptr(), flat.numel(), codes.data_ptr(), codebooks.data_ptr(),
                                   scales.data_ptr(), out.data_ptr(), g * d, n, g * d, d, nbits, m,
                                   _DTYPES[codebooks.dtype], torch.cuda.current_stream().cuda_stream)
    _lib.check(rc, "pq::embedding")
    return out.reshape(ids.shape + (g * d,))


def _gemm_cuda(input, codes, codebooks, scales, codebook_partition_sizes, bias):
    return _gemm_impl(input, codes, codebooks, scales, codebook_partition_sizes, bias, _lib.ALGO_AUTO)


def _gemm_algo_cuda(input, codes, codebooks, scales, codebook_partition_sizes, bias, algo):
    return _gemm_impl(input, codes, codebooks, scales, codebook_partition_sizes, bias, algo)


def _gemm_silu_mul_cuda(input, codes, codebooks, scales, codebook_partition_sizes, bias, algo):
    return _gemm_impl(input, codes, codebooks, scales, codebook_partition_sizes, bias, algo, silu_mul=True)


def _silu_and_mul_cuda(input):
    _require_cuda(input)
    if input.dtype not in _DTYPES or input.shape[-1] % 4:
        raise RuntimeError("PQ: silu_and_mul takes fp16/bf16 [..., 2k] with k even")
    k = input.shape[-1] // 2
    x = input.reshape(-1, 2 * k)
    if x.stride(-1) != 1 or x.stride(0) % 2 or x.data_ptr() % 4:
        x = x.contiguous()
    out = torch.empty((x.shape[0], k), dtype=input.dtype, device=input.device)
    if x.shape[0]:
        with torch.cuda.device(input.device):
            rc = _lib.load().pq_b200_silu_and_mul(x.data_ptr(), x.stride(0), out.data_ptr(), k, x.shape[0], k,
                                                  _DTYPES[input.dtype], torch.cuda.current_stream().cuda_stream)
        _lib.check(rc, "pq::silu_and_mul")
    return out.reshape(input.shape[:-1] + (k,))


def _dequant_cuda(codes, codebooks, codebook_partition_sizes):
    return _dequant_impl(codes, codebooks, None, codebook_partition_sizes)


def _dequant_scaled_cuda(codes, codebooks, scales, codebook_partition_sizes):
    return _dequant_impl(codes, codebooks, scales, codebook_partition_sizes)


_LIBRARY.impl("gemm", _gemm_cuda, "CUDA")
_LIBRARY.impl("gemm_algo", _gemm_algo_cuda, "CUDA")
_LIBRARY.impl("gemm_silu_mul", _gemm_silu_mul_cuda, "CUDA")
_LIBRARY.impl("silu_and_mul", _silu_and_mul_cuda, "CUDA")
_LIBRARY.impl("gemm_fused", _gemm_fused_cuda, "CUDA")
_LIBRARY.impl("fused_add_rms_norm", _fused_add_rms_norm_cuda, "CUDA")
_LIBRARY.impl("embedding", _embedding_cuda, "CUDA")
_LIBRARY.impl("dequant", _dequant_cuda, "CUDA")
_LIBRARY.impl("dequant_scaled", _dequant_scaled_cuda, "CUDA")


# ---- fake (meta) implementations: shape rules of vllm/_custom_ops.py:347-371 -------
def _gemm_fake(input, codes, codebooks, scales, codebook_partition_sizes, bias):
    out_features = codes.size(0) * codebooks.size(2)
    flat_input = input.reshape((-1, input.size(-1)))
    flat_output = torch.empty((flat_input.size(0), out_features), dtype=input.dtype,
                              device=input.device)
    output_sizes = list(input.shape)
    output_sizes.pop()
    output_sizes.append(-1)
    return flat_output.reshape(tuple(output_sizes))


def _gemm_algo_fake(input, codes, codebooks, scales, codebook_partition_sizes, bias, algo):
    return _gemm_fake(input, codes, codebooks, scales, codebook_partition_sizes, bias)


def _dequant_fake(codes, codebooks, codebook_partition_sizes):
    in_features = codes.size(1) * codebooks.size(3)
    out_features = codes.size(0)
    return torch.empty((out_features, in_features), dtype=codebooks.dtype, device=codebooks.device)


def _dequant_scaled_fake(codes, codebooks, scales, codebook_partition_sizes):
    return _dequant_fake(codes, codebooks, codebook_partition_sizes)


torch.library.register_fake("pq::gemm")(_gemm_fake)
torch.library.register_fake("pq::gemm_algo")(_gemm_algo_fake)
torch.library.register_fake("pq::gemm_silu_mul")(_gemm_algo_fake)
torch.library.register_fake("pq::silu_and_mul")(
    lambda input: torch.empty(input.shape[:-1] + (input.shape[-1] // 2,), dtype=input.dtype, device=input.device))
def _gemm_fused_fake(input, codes, codebooks, scales, codebook_partition_sizes, bias, residual, norm_ssq,
                     norm_weight, norm_eps, allreduce, silu_mul, want_ssq, codes_padded=None):
    n = codes.size(0)
    y = torch.empty(input.shape[:-1] + (n,), dtype=input.dtype, device=input.device)
    t = y.numel() // n
    # the slot count is the kernel's CTA count (pq_b200_gemm_fused_slots); a symbolic stand-in here
    return y, torch.empty((t, 1) if want_ssq else (0,), dtype=torch.float32, device=input.device)


torch.library.register_fake("pq::gemm_fused")(_gemm_fused_fake)
torch.library.register_fake("pq::fused_add_rms_norm")(lambda input, residual, weight, epsilon: None)
torch.library.register_fake("pq::embedding")(
    lambda ids, codes, codebooks, scales: torch.empty(ids.shape + (codes.size(1) * codebooks.size(3),),
                                                      dtype=codebooks.dtype, device=codebooks.device))
torch.library.register_fake("pq::dequant")(_dequant_fake)
torch.library.register_fake("pq::dequant_scaled")(_dequant_scaled_fake)


# ---- python wrappers, named after vllm/_custom_ops.py:646-658 ------------------------
def pq_gemm(input: torch.Tensor, codes: torch.Tensor, codebooks: torch.Tensor,
            scales: torch.Tensor, codebook_partition_sizes: List[int],
            bias: Optional[torch.Tensor], algo: int = _lib.ALGO_AUTO) -> torch.Tensor:
    if algo == _lib.ALGO_AUTO:
        return torch.ops.pq.gemm(input, codes, codebooks, scales, codebook_partition_sizes, bias)
    return torch.ops.pq.gemm_algo(input, codes, codebooks, scales, codebook_partition_sizes, bias, algo)


def pq_gemm_silu_mul(input: torch.Tensor, codes: torch.Tensor, codebooks: torch.Tensor,
                     scales: torch.Tensor, codebook_partition_sizes: List[int],
                     bias: Optional[torch.Tensor], algo: int = _lib.ALGO_AUTO) -> torch.Tensor:
    """pq_gemm(SiLU(input[..., :k]) * input[..., k:], ...): the MLP's act_fn + down_proj
    (llama.py:89-93) in one call; the decode table GEMV applies the activation on its x load."""
    return torch.ops.pq.gemm_silu_mul(input, codes, codebooks, scales, codebook_partition_sizes, bias, algo)


def silu_and_mul(input: torch.Tensor) -> torch.Tensor:
    """SiluAndMul.forward_cuda (layers/activation.py:65-84) on this library's kernel."""
    return torch.ops.pq.silu_and_mul(input)


def pq_gemm_fused(input: torch.Tensor, codes: torch.Tensor, codebooks: torch.Tensor, scales: torch.Tensor,
                  codebook_partition_sizes: List[int], bias: Optional[torch.Tensor] = None, *,
                  residual: Optional[torch.Tensor] = None, norm_ssq: Optional[torch.Tensor] = None,
                  norm_weight: Optional[torch.Tensor] = None, norm_eps: float = 1e-6, allreduce: bool = False,
                  silu_mul: bool = False, want_ssq: bool = False, codes_padded: Optional[torch.Tensor] = None):
    """(y, ssq): the decode table GEMV with its fused prologue / epilogue (include/pq_b200.h,
    pq_b200_gemm_fused).  Replaces apply -> all_reduce -> fused_add_rms_norm -> ... of
    linear.py:1209-1213 / llama.py:272-286 at decode."""
    return torch.ops.pq.gemm_fused(input, codes, codebooks, scales, codebook_partition_sizes, bias, residual,
                                   norm_ssq, norm_weight, norm_eps, allreduce, silu_mul, want_ssq, codes_padded)


def fused_add_rms_norm(input: torch.Tensor, residual: Optional[torch.Tensor], weight: torch.Tensor,
                       epsilon: float) -> None:
    """In place, as the reference's ops.fused_add_rms_norm / rms_norm (vllm/_custom_ops.py,
    csrc/layernorm_kernels.cu:17-98): residual += input; input = RMSNorm(residual) * weight."""
    torch.ops.pq.fused_add_rms_norm(input, residual, weight, epsilon)


def pq_embedding(ids: torch.Tensor, codes: torch.Tensor, codebooks: torch.Tensor, scales: torch.Tensor) -> torch.Tensor:
    """F.embedding(ids, W) over the PQ weight W (one output partition): rows of the scaled dequantised
    weight, bit-identical to pq_dequant(..., scales)[ids]; ids outside [0, N) give zero rows."""
    return torch.ops.pq.embedding(ids, codes, codebooks, scales)


def pq_dequant(codes: torch.Tensor, codebooks: torch.Tensor,
               codebook_partition_sizes: List[int],
               scales: Optional[torch.Tensor] = None) -> torch.Tensor:
    if scales is None:
        return torch.ops.pq.dequant(codes, codebooks, codebook_partition_sizes)
    return torch.ops.pq.dequant_scaled(codes, codebooks, scales, codebook_partition_sizes)


def set_pdl(enabled: bool) -> None:
    """Programmatic dependent launch for the decode kernels (see csrc/pq_common.cuh)."""
    _lib.load().pq_b200_set_pdl(1 if enabled else 0)
