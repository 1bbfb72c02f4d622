"""PQConfig / PQLinearMethod: the product-quantization linear method.

Drop-in for the reference's AQLMConfig / AQLMLinearMethod
(vllm/model_executor/layers/quantization/aqlm.py:159-374): same constructor
arguments, same HF ``quantization_config`` keys, same parameter names, shapes,
dtypes and loader attributes in ``create_weights`` (aqlm.py:253-323), same
``apply(layer, x, bias)`` contract - so the reference's unmodified weight
loaders (linear.py:370-408, 494-640, 873-1058, 1144-1180) fill the parameters.

What changes is below the op boundary: every (nbits, d, M) format runs on the
sm_100a kernels in csrc/ (the reference has CUDA kernels for 1x16 and 2x8 only
and sends everything else through torch ops, aqlm.py:344-374), fp16 *and* bf16
are accepted (reference: fp16 only, aqlm.py:193-194,238-239), and there is no
CPU path.
"""
from __future__ import annotations

from typing import Any, Dict, List, Optional

import torch
from torch.nn.parameter import Parameter

from . import _custom_ops as ops
from .plugin_api import (LinearMethodBase, QuantizationConfig, register_quantization_config,
                         set_weight_attrs)


def _from_keys(config: Dict[str, Any], key: str) -> Any:
    """base_config.py:104-112 (get_from_keys), same error."""
    if key in config:
        return config[key]
    raise ValueError(f"Cannot find any of {[key]} in the model's quantization config.")


def get_int_dtype(nbits: int) -> torch.dtype:
    """Storage type of a code (aqlm.py:20-29); a 4-bit code still takes an int8."""
    for limit, dt in ((8, torch.int8), (16, torch.int16), (32, torch.int32), (64, torch.int64)):
        if nbits <= limit:
            return dt
    raise ValueError(f"No dtype available for {nbits}-bit codebooks")


class PQConfigBody:
    """Method bodies of the PQ config (aqlm.py:159-217), independent of the ABC they are mounted on:
    ``make_classes`` combines them with this package's restated ABCs (stand-alone) or with vLLM's own
    ``QuantizationConfig`` (``vllm_plugin.register``).  No zero-argument ``super()`` in here - its
    ``__class__`` cell would pin the bodies to one base."""

    _config_base: type = object                              # set by make_classes
    _method_cls: type = None                                 # "
    _is_linear = staticmethod(lambda layer: False)           # "
    _is_vocab = staticmethod(lambda layer: False)            # ": VocabParallelEmbedding / ParallelLMHead

    def __init__(self, in_group_size: int, nbits_per_codebook: int, num_codebooks: int,
                 out_group_size: int = 1, lm_head_quantized: bool = False) -> None:
        type(self)._config_base.__init__(self)
        # The reference leaves embeddings and lm_head dense (aqlm.py:213-217 returns None for anything
        # but a LinearBase).  With "lm_head_quantized": true in the HF quantization_config (the key
        # vLLM's GPTQ / AWQ-marlin configs use for the same switch) the vocabulary layers take PQ
        # weights too: SURVEY.md 8f N4.
        self.lm_head_quantized = bool(lm_head_quantized)
        self.in_group_size = in_group_size
        self.nbits_per_codebook = nbits_per_codebook
        self.num_codebooks = num_codebooks
        self.out_group_size = out_group_size
        assert self.out_group_size == 1                      # aqlm.py:178-179
        if nbits_per_codebook > 16:
            raise ValueError("PQ kernels support at most 16 bits per code")
        self.pack_factor = self.in_group_size * self.out_group_size

    def __repr__(self) -> str:
        return (f"PQConfig(in_group_size={self.in_group_size}, "
                f"nbits_per_codebook={self.nbits_per_codebook}, "
                f"num_codebooks={self.num_codebooks}, "
                f"out_group_size={self.out_group_size})")

    @classmethod
    def get_name(cls) -> str:
        return cls._registered_name

    @classmethod
    def get_supported_act_dtypes(cls) -> List[torch.dtype]:
        return [torch.half, torch.bfloat16]

    @classmethod
    def get_min_capability(cls) -> int:
        return 100                                           # sm_100a only

    @classmethod
    def get_config_filenames(cls) -> List[str]:
        return []

    @classmethod
    def from_config(cls, config: Dict[str, Any]) -> "PQConfig":
        return cls(_from_keys(config, "in_group_size"), _from_keys(config, "nbits_per_codebook"),
                   _from_keys(config, "num_codebooks"), _from_keys(config, "out_group_size"),
                   bool(config.get("lm_head_quantized", False)))

    @classmethod
    def override_quantization_method(cls, hf_quant_cfg, user_quant) -> Optional[str]:
        # An AQLM checkpoint can be served by this method when the user asks for it.
        name = cls._registered_name
        if user_quant == name and hf_quant_cfg.get("quant_method") in ("aqlm", name):
            return name
        return None

    def get_quant_method(self, layer: torch.nn.Module, prefix: str = "") -> Optional["PQLinearMethod"]:
        if type(self)._is_linear(layer):
            return type(self)._method_cls(self)
        if self.lm_head_quantized and type(self)._is_vocab(layer):
            return type(self)._method_cls(self)              # PQ lm_head / embedding (N4)
        return None                                          # embeddings / lm_head stay dense


class PQLinearMethodBody:
    """Method bodies of the PQ linear method (aqlm.py:220-374); see PQConfigBody."""

    def __init__(self, quant_config: "PQConfig"):
        self.quant_config = quant_config

    def create_weights(self, layer: torch.nn.Module, input_size_per_partition: int,
                       output_partition_sizes: List[int], input_size: int, output_size: int,
                       params_dtype: torch.dtype, **extra_weight_attrs):
        del output_size, input_size
        cfg = self.quant_config
        if params_dtype not in (torch.half, torch.bfloat16):
            raise ValueError("Only half and bfloat16 are currently supported by pq")
        if input_size_per_partition % cfg.in_group_size != 0:
            raise ValueError("The input size is not aligned with the quantized weight shape. "
                             "This can be caused by too large tensor parallel size.")
        output_size_per_partition = sum(output_partition_sizes)
        if output_size_per_partition % cfg.out_group_size != 0:
            raise ValueError("The output size is not aligned with the quantized weight shape. "
                             "This can be caused by too large tensor parallel size.")

        codes = Parameter(torch.empty(output_size_per_partition,
                                      input_size_per_partition // cfg.pack_factor,
                                      cfg.num_codebooks,
                                      dtype=get_int_dtype(cfg.nbits_per_codebook)),
                          requires_grad=False)
        set_weight_attrs(codes, {"input_dim": 1, "output_dim": 0, "packed_dim": 1,
                                 "pack_factor": cfg.pack_factor})

        codebooks = Parameter(torch.empty(cfg.num_codebooks * len(output_partition_sizes),
                                          2**cfg.nbits_per_codebook, cfg.out_group_size,
                                          cfg.in_group_size, dtype=params_dtype),
                              requires_grad=False)
        # metadata: fixed size, concatenated along dim 0 by the fused-layer loaders
        set_weight_attrs(codebooks, {"is_metadata": True,
                                     "output_partition_sizes": output_partition_sizes})

        scales = Parameter(torch.empty((output_size_per_partition // cfg.out_group_size, 1, 1, 1),
                                       dtype=params_dtype),
                           requires_grad=False)
        set_weight_attrs(scales, {"output_dim": 0, "packed_dim": 0,
                                  "pack_factor": cfg.out_group_size})

        for name, p in (("codes", codes), ("codebooks", codebooks), ("scales", scales)):
            layer.register_parameter(name, p)
            set_weight_attrs(p, extra_weight_attrs)

    def process_weights_after_loading(self, layer: torch.nn.Module) -> None:
        """The hook vLLM calls once after the checkpoint tensors are in place (base_config.py:37-42,
        model_loader/loader.py:157-168).  The kernels read the checkpoint layout directly, so the
        parameters stay as loaded (contiguity made explicit once, not per call).  One private repack
        (SURVEY.md 8f N3): 8-bit-or-narrower codes whose row is not a multiple of 16 bytes - the
        row-parallel shards of down_proj at TP = 4 / 8 (344 / 172 bytes) - get a copy with the row
        pitch padded to 16 bytes, ``layer.codes_padded`` [N, pitch]: TMA cannot address the unpadded
        rows, and the cp.async loader that serves them reads 1.27 x the bytes.  The decode table GEMV
        streams the copy; every other kernel (and pq::dequant) keeps reading ``codes``."""
        for name in ("codes", "codebooks", "scales"):
            p = getattr(layer, name)
            if not p.data.is_contiguous():
                p.data = p.data.contiguous()
        codes = layer.codes.data
        n, g, m = codes.shape
        row = g * m * codes.element_size()
        if codes.element_size() == 1 and row % 16 and row >= 64 and codes.is_cuda:
            pitch = (row + 15) // 16 * 16
            padded = torch.zeros((n, pitch), dtype=codes.dtype, device=codes.device)
            padded[:, :row] = codes.view(n, row)
            layer.codes_padded = padded
        elif hasattr(layer, "codes_padded"):
            del layer.codes_padded
        self._pad_k(layer)
        self._repack_nibbles(layer)

    @staticmethod
    def _k_padded_groups(g: int, m: int, d: int, elem_bytes: int) -> int:
        """Smallest G' >= G whose K' = G' * d is whole 64-wide k blocks (what the mma.sync and tcgen05
        kernels tile K by) and whose code row G' * M * elem_bytes is a multiple of 16 bytes."""
        gp = g
        while (gp * d) % 64 or (gp * m * elem_bytes) % 16:
            gp += 1
        return gp

    @classmethod
    def _pad_k(cls, layer: torch.nn.Module, force: bool = False) -> None:
        """K that is not a multiple of 64 - Llama-2-7B down_proj at TP = 8: K = 1376 - is outside the
        mma.sync and tcgen05 kernels, which leaves the scalar gather GEMV for every batch above the
        table GEMV's eight tokens (the only batches this copy is used for).  ``layer.codes_kpad`` [N, G', M] is the code tensor with zero codes
        appended up to K' = G' * d; ``_gemm_checkpoint`` feeds it an x zero-padded to K', so the extra
        groups add  <codeword[0], 0> = 0  exactly and the kernels see an ordinary K' problem.  For 8-bit
        codes this is the SAME storage as ``codes_padded`` (a view) whenever the two pitches agree."""
        if hasattr(layer, "codes_kpad"):
            del layer.codes_kpad
        codes = layer.codes.data
        n, g, m = codes.shape
        d = layer.codebooks.shape[3]
        if (g * d) % 64 == 0 or not (codes.is_cuda or force) or codes.element_size() > 2:
            return
        gp = cls._k_padded_groups(g, m, d, codes.element_size())
        padded = getattr(layer, "codes_padded", None)
        if padded is not None and padded.shape[1] == gp * m:
            layer.codes_kpad = padded.view(n, gp, m)
        else:
            kpad = torch.zeros((n, gp, m), dtype=codes.dtype, device=codes.device)
            kpad[:, :g] = codes
            layer.codes_kpad = kpad

    def _gemm_checkpoint(self, layer: torch.nn.Module, x: torch.Tensor, bias: Optional[torch.Tensor],
                         silu_mul: bool = False) -> torch.Tensor:
        """``pq::gemm`` (or ``pq::gemm_silu_mul``) over the checkpoint tensors - every token count and
        format, the C side picks the kernel - or, where ``_pad_k`` built one, over the K-padded copy."""
        codebooks = layer.codebooks
        parts = list(getattr(codebooks, "output_partition_sizes", []))
        kpad = getattr(layer, "codes_kpad", None)
        # above the table GEMV's eight tokens only: at decode the checkpoint layout already has its kernels
        # (the table GEMV's cp.async / padded-pitch loaders, the gather GEMV)
        if kpad is not None and x.numel() // max(x.shape[-1], 1) > 8:
            if silu_mul:
                x = ops.silu_and_mul(x)
            k_pad = kpad.shape[1] * codebooks.shape[3]
            x = torch.nn.functional.pad(x, (0, k_pad - x.shape[-1]))
            return ops.pq_gemm(x, kpad, codebooks, layer.scales, parts, bias)
        if silu_mul:
            return ops.pq_gemm_silu_mul(x, layer.codes, codebooks, layer.scales, parts, bias)
        return ops.pq_gemm(x, layer.codes, codebooks, layer.scales, parts, bias)

    @staticmethod
    def _repack_nibbles(layer: torch.nn.Module) -> None:
        """4-bit codes occupy a full int8 each in the checkpoint (aqlm.py:20-29: no sub-byte packing).
        Two of them packed into one byte index a 256-entry PRODUCT codebook, which is again a PQ
        layer the table GEMV already serves - half the bytes streamed, half the lookups:
          * M = 1: the codes of two consecutive groups -> one group of 2d elements whose codeword is the
            concatenation  cb'[lo | hi << 4] = (cb[lo] || cb[hi])            (d in {4, 8}, G even);
          * M = 2: the two codes of one group -> one code whose codeword is the sum
            cb'[lo | hi << 4] = cb0[lo] + cb1[hi], fp32 sum rounded once - exactly the dequantised
            weight of aqlm.py:60-65, so the GEMV sees the weights ``pq::dequant`` returns.
        Private copies (``codes_nibble`` [N, G', 1], ``codebooks_pair`` [P, 256, 1, d']), decode only
        (1..8 tokens); the checkpoint tensors stay for every other kernel and for ``pq::dequant``."""
        for name in ("codes_nibble", "codebooks_pair"):
            if hasattr(layer, name):
                delattr(layer, name)
        codes, cb = layer.codes.data, layer.codebooks.data
        n, g, m = codes.shape
        pm, e, _, d = cb.shape
        if e != 16 or not codes.is_cuda or codes.dtype != torch.int8:
            return
        idx = (codes.to(torch.int16) & 15)
        p = pm // m
        if m == 1 and d in (4, 8) and g % 2 == 0 and (g // 2) >= 64:
            packed = (idx[:, 0::2, 0] | (idx[:, 1::2, 0] << 4))
            lo = torch.arange(256, device=cb.device) & 15
            hi = torch.arange(256, device=cb.device) >> 4
            pair = torch.cat([cb[:, lo, 0, :], cb[:, hi, 0, :]], dim=-1)              # [P, 256, 2d]
        elif m == 2 and d in (4, 8, 16) and g >= 64:
            packed = (idx[:, :, 0] | (idx[:, :, 1] << 4))
            lo = torch.arange(256, device=cb.device) & 15
            hi = torch.arange(256, device=cb.device) >> 4
            cbp = cb.view(p, 2, 16, d)
            pair = (cbp[:, 0, lo, :].float() + cbp[:, 1, hi, :].float()).to(cb.dtype)   # [P, 256, d]
        else:
            return
        packed = packed.to(torch.uint8).view(torch.int8).unsqueeze(-1).contiguous()     # signed storage, as codes
        layer.codes_nibble = packed
        layer.codebooks_pair = pair.unsqueeze(2).contiguous()

    @staticmethod
    def _decode_tensors(layer: torch.nn.Module):
        """(codes, codebooks) the decode table GEMV streams: the nibble-packed private copy of a 4-bit
        layer where there is one (process_weights_after_loading), else the checkpoint tensors."""
        nib = getattr(layer, "codes_nibble", None)
        if nib is not None:
            return nib, layer.codebooks_pair
        return layer.codes, layer.codebooks

    def _decode_fused_ok(self, layer: torch.nn.Module, x: torch.Tensor) -> bool:
        codes, codebooks = self._decode_tensors(layer)
        n, g, m = codes.shape
        d, nbits = codebooks.shape[3], codebooks.shape[1].bit_length() - 1
        t = x.numel() // x.shape[-1] if x.shape[-1] else 0
        return x.is_cuda and t >= 1 and ops.gemm_fused_supported(t, n, g * d, d, nbits, m)

    def apply(self, layer: torch.nn.Module, x: torch.Tensor,
              bias: Optional[torch.Tensor] = None) -> torch.Tensor:
        codebooks = layer.codebooks
        output_partition_sizes = getattr(codebooks, "output_partition_sizes", [])
        padded = getattr(layer, "codes_padded", None)
        nibble = getattr(layer, "codes_nibble", None)
        if (padded is not None or nibble is not None) and self._decode_fused_ok(layer, x):
            # decode over a private copy (process_weights_after_loading): padded rows / packed nibbles
            dc, dcb = self._decode_tensors(layer)
            return ops.pq_gemm_fused(x, dc, dcb, layer.scales, list(output_partition_sizes), bias,
                                     codes_padded=padded if nibble is None else None)[0]
        # One op for every token count and format: the C side picks GEMV (decode) or the
        # tcgen05 GEMM (prefill). Reference dispatch for comparison: aqlm.py:344-374.
        return self._gemm_checkpoint(layer, x, bias)

    def apply_silu_mul(self, layer: torch.nn.Module, x: torch.Tensor,
                       bias: Optional[torch.Tensor] = None) -> torch.Tensor:
        """apply(SiLU(x[..., :k]) * x[..., k:]): act_fn + down_proj of the MLP (llama.py:89-93)."""
        codebooks = layer.codebooks
        parts = list(getattr(codebooks, "output_partition_sizes", []))
        padded = getattr(layer, "codes_padded", None)
        nibble = getattr(layer, "codes_nibble", None)
        if (padded is not None or nibble is not None) and self._decode_fused_ok(layer, x[..., : x.shape[-1] // 2]):
            dc, dcb = self._decode_tensors(layer)
            return ops.pq_gemm_fused(x, dc, dcb, layer.scales, parts, bias, silu_mul=True,
                                     codes_padded=padded if nibble is None else None)[0]
        return self._gemm_checkpoint(layer, x, bias, silu_mul=True)

    def apply_fused(self, layer: torch.nn.Module, x: torch.Tensor, bias: Optional[torch.Tensor] = None, *,
                    residual: Optional[torch.Tensor] = None, norm=None, allreduce: bool = False,
                    silu_mul: bool = False, want_ssq: bool = False):
        """(z, ssq) with  z = all_reduce(apply(act(norm(x)))) + residual  - the chain the reference runs
        as separate launches around a row-parallel layer (linear.py:1209-1213, llama.py:272-286,
        layernorm_kernels.cu:52-98).  ``norm`` = (ssq or None, weight, eps): RMSNorm of x, its sum of
        squares coming from the producer's ``ssq`` ([T, slots] fp32) when there is one.  At decode
        (1..8 tokens, nbits <= 8) everything rides on the table GEMV's x load and epilogue; otherwise
        the same steps run as separate launches with the same roundings."""
        from . import distributed as D
        codes, codebooks = layer.codes, layer.codebooks
        parts = list(getattr(codebooks, "output_partition_sizes", []))
        dc, dcb = self._decode_tensors(layer)        # what the table GEMV would stream (nibble-packed copy?)
        n, g, m = dc.shape
        d, nbits = dcb.shape[3], dcb.shape[1].bit_length() - 1
        t = x.numel() // x.shape[-1] if x.shape[-1] else 0
        comm = D.fused_allreduce_comm() if allreduce else None
        fused = (x.is_cuda and t >= 1 and ops.gemm_fused_supported(t, n, g * d, d, nbits, m)
                 and (not allreduce or (comm is not None and t * n <= comm.LL_SLOT_ELEMS))
                 and not (norm is not None and silu_mul))
        if norm is not None and (not fused or norm[0] is None):
            x = self._rms_norm(x, *norm)
            norm = None
        if fused:
            ns, nw, ne = norm if norm is not None else (None, None, 1e-6)
            y, ssq = ops.pq_gemm_fused(x, dc, dcb, layer.scales, parts, bias, residual=residual,
                                       norm_ssq=ns, norm_weight=nw, norm_eps=ne, allreduce=allreduce,
                                       silu_mul=silu_mul, want_ssq=want_ssq,
                                       codes_padded=getattr(layer, "codes_padded", None) if dc is codes else None)
            return y, (ssq if want_ssq else None)
        y = self._gemm_checkpoint(layer, x, bias, silu_mul=silu_mul)
        if allreduce:
            y = D.tensor_model_parallel_all_reduce(y)
        if residual is not None:
            y = y + residual.reshape(y.shape)
        ssq = y.float().pow(2).sum(-1).reshape(-1, 1) if want_ssq else None
        return y, ssq

    @staticmethod
    def _rms_norm(x: torch.Tensor, ssq: Optional[torch.Tensor], weight: torch.Tensor, eps: float) -> torch.Tensor:
        """RMSNorm(x) * weight as its own step; roundings of layernorm_kernels.cu:91-97."""
        if ssq is None:
            out = x.clone()
            ops.fused_add_rms_norm(out, None, weight, eps)
            return out
        rstd = torch.rsqrt(ssq.reshape(-1, ssq.shape[-1]).sum(-1, keepdim=True) / x.shape[-1] + eps)
        return ((x.reshape(-1, x.shape[-1]).float() * rstd).to(x.dtype) * weight).reshape(x.shape)

    def embedding(self, layer: torch.nn.Module, input_: torch.Tensor) -> torch.Tensor:
        """Rows ``input_`` of the scaled dequantised weight (QuantizeMethodBase.embedding,
        base_config.py:30-35; UnquantizedEmbeddingMethod.embedding is F.embedding over the weight,
        vocab_parallel_embedding.py:41-43).  Bit-identical to ``dequantize(layer)[input_]``."""
        return ops.pq_embedding(input_, layer.codes, layer.codebooks, layer.scales)

    def dequantize(self, layer: torch.nn.Module, scaled: bool = True) -> torch.Tensor:
        """Dense [N, K] weight (aqlm.py:37-74); bit-exact with the reference."""
        codebooks = layer.codebooks
        parts = list(getattr(codebooks, "output_partition_sizes", []))
        return ops.pq_dequant(layer.codes, codebooks, parts, layer.scales if scaled else None)


def make_classes(config_base: type, method_base: type, is_linear, name: str = "pq", is_vocab=None):
    """(PQConfig, PQLinearMethod) as real subclasses of the given ABCs.  ``is_linear(layer)`` is the
    ``isinstance(layer, LinearBase)`` test of aqlm.py:213-217 for whichever LinearBase is in use."""
    method_cls = type("PQLinearMethod", (PQLinearMethodBody, method_base),
                      {"__doc__": PQLinearMethodBody.__doc__, "__module__": __name__})
    config_cls = type("PQConfig", (PQConfigBody, config_base), {
        "__doc__": "Config class for PQ (AQLM-format) checkpoints.  aqlm.py:159-217.",
        "__module__": __name__,
        "_config_base": config_base, "_method_cls": method_cls, "_is_linear": staticmethod(is_linear),
        "_is_vocab": staticmethod(is_vocab if is_vocab is not None else (lambda layer: False)),
        "_registered_name": name,
    })
    return config_cls, method_cls


def _is_own_linear(layer) -> bool:
    from .linear import is_linear_layer
    return is_linear_layer(layer)


def _is_own_vocab(layer) -> bool:
    from .embedding import VocabParallelEmbedding
    return isinstance(layer, VocabParallelEmbedding)


PQConfig, PQLinearMethod = make_classes(QuantizationConfig, LinearMethodBase, _is_own_linear, is_vocab=_is_own_vocab)
PQConfig = register_quantization_config("pq")(PQConfig)
