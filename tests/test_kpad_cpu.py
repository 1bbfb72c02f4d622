"""K that is not whole 64-wide k blocks (Llama-2-7B down_proj at TP = 8: K = 1376): the private K-padded
copy of process_weights_after_loading and the routing of apply / apply_silu_mul / apply_fused over it.
Host logic only - ``pq::gemm`` is replaced by the oracle's fp64 contraction, in which zero products
change nothing, so the padded and the unpadded problem must give the same bits."""
import pytest
import torch

import vllm_pq_b200 as P
from oracle import pq_oracle as O
from vllm_pq_b200 import _custom_ops as ops
from vllm_pq_b200 import _lib


@pytest.fixture
def cpu_ops(monkeypatch):
    calls = []

    def cpu_gemm(x, codes, codebooks, scales, parts, bias, algo=0):
        calls.append(("gemm", tuple(x.shape), tuple(codes.shape)))
        return O.linear_fp32(x, codes, codebooks, scales, parts or [codes.shape[0]], bias).to(x.dtype)

    def cpu_gemm_silu(x, codes, codebooks, scales, parts, bias, algo=0):
        calls.append(("gemm_silu_mul", tuple(x.shape), tuple(codes.shape)))
        return cpu_gemm(O.silu_and_mul(x), codes, codebooks, scales, parts, bias)

    monkeypatch.setattr(ops, "pq_gemm", cpu_gemm)
    monkeypatch.setattr(ops, "pq_gemm_silu_mul", cpu_gemm_silu)
    monkeypatch.setattr(ops, "silu_and_mul", O.silu_and_mul)
    return calls


def make_layer(k, fmt, n=256):
    b, d, m = fmt
    c = O.make_case([n], k, b, d, m, t=12, dtype=torch.float16, seed=5, bias=True)
    c["codebooks"] = (c["codebooks"].float() / (k * m) ** 0.5).half()
    lin = P.RowParallelLinear(k, n, bias=True, params_dtype=torch.float16, quant_config=P.PQConfig(d, b, m, 1))
    for key in ("codes", "codebooks", "scales", "bias"):
        getattr(lin, key).data.copy_(c[key])
    return lin, c


@pytest.mark.parametrize("fmt", [(8, 8, 1), (8, 8, 2), (16, 8, 1), (4, 4, 1), (12, 16, 1)])
def test_k_padded_copy_and_routing(cpu_ops, fmt):
    b, d, m = fmt
    k = 1376
    lin, c = make_layer(k, fmt)
    qm = lin.quant_method
    qm.process_weights_after_loading(lin)
    assert not hasattr(lin, "codes_kpad")                  # CPU tensors: no private copies are built
    y_plain, _ = lin(c["x"])
    qm._pad_k(lin, force=True)
    g = k // d
    gp = qm._k_padded_groups(g, m, d, lin.codes.element_size())
    assert (gp * d) % 64 == 0 and (gp * m * lin.codes.element_size()) % 16 == 0 and gp - g < 64
    assert lin.codes_kpad.shape == (256, gp, m) and lin.codes_kpad.dtype == lin.codes.dtype
    assert torch.equal(lin.codes_kpad[:, :g], lin.codes.data) and not lin.codes_kpad[:, g:].any()
    # the padded problem is one the mma.sync / tcgen05 kernels take, the unpadded one is not
    lib = _lib.load()
    if d == 8:
        assert not lib.pq_b200_algo_supported(_lib.ALGO_GEMM_TC, 64, 4096, k, d, b, m)
        assert lib.pq_b200_algo_supported(_lib.ALGO_GEMM_TC, 64, 4096, gp * d, d, b, m)
        assert lib.pq_b200_pick_algo(64, 4096, gp * d, d, b, m) == _lib.ALGO_GEMM_TC
    del cpu_ops[:]
    y, _ = lin(c["x"])
    assert cpu_ops == [("gemm", (12, gp * d), (256, gp, m))]
    assert torch.equal(y, y_plain)
    # SiLU*mul input: the activation first, then the padded linear
    gu = torch.randn(12, 2 * k, generator=torch.Generator().manual_seed(1)).half()
    del lin.codes_kpad
    want, _ = lin(gu, input_act="silu_and_mul")
    qm._pad_k(lin, force=True)
    del cpu_ops[:]
    got, _ = lin(gu, input_act="silu_and_mul")
    assert cpu_ops == [("gemm", (12, gp * d), (256, gp, m))] and torch.equal(got, want)
    # the separate-launch form of the fused chain (residual add, sum of squares) over the same copy
    res = torch.randn(12, 256, generator=torch.Generator().manual_seed(2)).half()
    out, _, ssq = lin(c["x"], residual=res, want_ssq=True)
    assert torch.equal(out, y_plain + res) and ssq.shape == (12, 1)
    # eight tokens or fewer keep the checkpoint layout (the decode kernels serve it)
    del cpu_ops[:]
    lin(c["x"][:8])
    assert cpu_ops == [("gemm", (8, k), (256, g, m))]
    # dequantize and the parameters are untouched
    assert dict(lin.named_parameters()).keys() == {"codes", "codebooks", "scales", "bias"}
    # a second call of the hook drops what no longer applies
    qm.process_weights_after_loading(lin)
    assert not hasattr(lin, "codes_kpad")


def test_no_copy_for_whole_k_blocks_and_shared_storage():
    lin, _ = make_layer(2752, (8, 8, 1))                   # TP = 4 shard: K = 43 * 64, pitch 344 (unaligned)
    lin.quant_method._pad_k(lin, force=True)
    assert not hasattr(lin, "codes_kpad")                  # the tcgen05 kernel takes K = 2752 as it is
    lin, _ = make_layer(1376, (8, 8, 1))
    n, g, m = lin.codes.shape
    lin.codes_padded = torch.zeros(n, 176, dtype=torch.int8)      # what the hook builds on the device
    lin.codes_padded[:, :g] = lin.codes.data.view(n, g)
    lin.quant_method._pad_k(lin, force=True)
    assert lin.codes_kpad.data_ptr() == lin.codes_padded.data_ptr()        # one copy serves both
    assert torch.equal(lin.codes_kpad[:, :g], lin.codes.data)
