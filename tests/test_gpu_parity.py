"""Parity of the CUDA path (through the torch op -> ctypes -> C ABI) against the oracle.
  P1/P2  dequantised weights: bit-exact                        
  P3     layer outputs: tolerance, stated below
"""
import pytest
import torch

import vllm_pq_b200 as P
from oracle import pq_oracle as O
from tests import golden_util as G
from vllm_pq_b200 import _lib

pytestmark = pytest.mark.gpu

# P3 gate: rms-relative error vs the exact (fp64) contraction
RMS_REL_TOL = {torch.float16: 2e-3, torch.bfloat16: 1.6e-2}


def dev(c):
    out = dict(c)
    for k in ("codes", "codebooks", "scales", "x", "bias"):
        if c.get(k) is not None:
            out[k] = c[k].cuda()
    return out


def check_output(y, c, algo_name=""):
    ref64 = O.linear_fp32(c["x"].cpu(), c["codes"].cpu(), c["codebooks"].cpu(), c["scales"].cpu(),
                          c["partition_sizes"], None if c["bias"] is None else c["bias"].cpu())
    err = O.error_report(y.cpu(), ref64)
    y_ref = O.generic_dequantize_gemm(c["x"].cpu(), c["codes"].cpu(), c["codebooks"].cpu(),
                                      c["scales"].cpu(), c["partition_sizes"],
                                      None if c["bias"] is None else c["bias"].cpu())
    err_ref = O.error_report(y_ref, ref64)
    dt = c["x"].dtype
    assert err["rms_rel"] <= RMS_REL_TOL[dt], (algo_name, err, err_ref)
    # no worse than 2x the reference's own low-precision path (plus one output ulp of slack)
    assert err["rms_rel"] <= 2.0 * err_ref["rms_rel"] + 1e-6, (algo_name, err, err_ref)
    assert err["rel_inf"] <= 2.0 * err_ref["rel_inf"] + 2 ** (-10 if dt == torch.float16 else -7), (algo_name, err, err_ref)
    return err, err_ref


@pytest.mark.parametrize("name", G.names())
def test_golden_dequant_bit_exact(name):
    c = dev(G.load(name))
    wu = P.ops.pq_dequant(c["codes"], c["codebooks"], c["partition_sizes"])
    ws = P.ops.pq_dequant(c["codes"], c["codebooks"], c["partition_sizes"], c["scales"])
    assert G.same_bits(wu, c["w_unscaled"])
    assert G.same_bits(ws, c["w_scaled"])


@pytest.mark.parametrize("name", G.names())
@pytest.mark.parametrize("algo", [_lib.ALGO_AUTO, _lib.ALGO_GEMV_DIRECT, _lib.ALGO_GEMV_LUT, _lib.ALGO_GEMM_TC,
                                  _lib.ALGO_GEMM_MMA, _lib.ALGO_GEMV_LUT_TMA])
def test_golden_gemm(name, algo):
    c = dev(G.load(name))
    lib = _lib.load()
    t, k, n = c["x"].shape[0], c["x"].shape[1], c["codes"].shape[0]
    if not lib.pq_b200_algo_supported(algo, t, n, k, c["d"], c["nbits"], c["m"]):
        pytest.skip(f"{_lib.ALGO_NAMES[algo]} kernel does not cover this format")
    y = P.ops.pq_gemm(c["x"], c["codes"], c["codebooks"], c["scales"], c["partition_sizes"], c["bias"], algo=algo)
    assert y.shape == c["y"].shape and y.dtype == c["y"].dtype
    check_output(y, c, _lib.ALGO_NAMES[algo])
    # and against what the reference itself produced for these inputs (golden y), loosely
    torch.testing.assert_close(y.cpu().float(), c["y"].float(), rtol=2e-2, atol=2e-2 * float(c["y"].float().abs().max()))


SWEEP = [  # (b, d, M)  - SURVEY.md 7.2 + BASELINE config 5 grid
    (8, 8, 1), (8, 8, 2), (16, 8, 1), (4, 4, 1), (4, 8, 1), (4, 16, 1), (8, 4, 1), (8, 16, 1),
    (12, 4, 1), (12, 8, 1), (12, 16, 1), (8, 8, 4), (8, 8, 3), (6, 8, 2), (16, 8, 2),
]


@pytest.mark.parametrize("fmt", SWEEP)
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
def test_dequant_bit_exact_sweep(fmt, dtype):
    b, d, m = fmt
    for parts, k in (([96, 32, 32], 64 * d), ([1], d), ([130], 43 * d)):
        c = O.make_case(parts, k, b, d, m, t=1, dtype=dtype, seed=11, positive_scales=False)
        g = dev(c)
        wu = P.ops.pq_dequant(g["codes"], g["codebooks"], parts)
        ws = P.ops.pq_dequant(g["codes"], g["codebooks"], parts, g["scales"])
        assert G.same_bits(wu, O.dequantize_partitions(c["codes"], c["codebooks"], parts, None)), (fmt, parts)
        assert G.same_bits(ws, O.dequantize_partitions(c["codes"], c["codebooks"], parts, c["scales"])), (fmt, parts)


@pytest.mark.parametrize("fmt", SWEEP)
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("t", [1, 2, 3, 5, 8])
def test_gemm_sweep(fmt, dtype, t):
    b, d, m = fmt
    parts, k = [128, 64, 72], 64 * d
    c = O.make_case(parts, k, b, d, m, t=t, dtype=dtype, seed=5 + t, bias=(t % 2 == 0))
    g = dev(c)
    y = P.ops.pq_gemm(g["x"], g["codes"], g["codebooks"], g["scales"], parts, g["bias"])
    check_output(y, c)
    yd = P.ops.pq_gemm(g["x"], g["codes"], g["codebooks"], g["scales"], parts, g["bias"], algo=_lib.ALGO_GEMV_DIRECT)
    check_output(yd, c, "direct")
    n = sum(parts)
    if _lib.load().pq_b200_algo_supported(_lib.ALGO_GEMV_LUT, t, n, k, d, b, m):
        yl = P.ops.pq_gemm(g["x"], g["codes"], g["codebooks"], g["scales"], parts, g["bias"], algo=_lib.ALGO_GEMV_LUT)
        check_output(yl, c, "lut")
        # split-K partials are summed in a fixed order: the result is reproducible bit for bit
        yl2 = P.ops.pq_gemm(g["x"], g["codes"], g["codebooks"], g["scales"], parts, g["bias"], algo=_lib.ALGO_GEMV_LUT)
        assert torch.equal(yl, yl2)
    if _lib.load().pq_b200_algo_supported(_lib.ALGO_GEMV_LUT_TMA, t, n, k, d, b, m):
        yt = P.ops.pq_gemm(g["x"], g["codes"], g["codebooks"], g["scales"], parts, g["bias"], algo=_lib.ALGO_GEMV_LUT_TMA)
        check_output(yt, c, "lut_tma")
        yt2 = P.ops.pq_gemm(g["x"], g["codes"], g["codebooks"], g["scales"], parts, g["bias"], algo=_lib.ALGO_GEMV_LUT_TMA)
        assert torch.equal(yt, yt2)


LUT_TMA_SWEEP = [(8, 8, 1), (8, 8, 2), (8, 8, 4), (4, 8, 1), (6, 8, 2), (8, 4, 1), (8, 16, 1), (5, 16, 2), (7, 4, 4)]


@pytest.mark.parametrize("fmt", LUT_TMA_SWEEP)
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("t", [1, 2, 5, 8])
@pytest.mark.parametrize("shape", [([128, 64, 72], 64), ([65], 80), ([1024, 48], 256), ([200], 96), ([2500, 700], 144),
                                   ([1024, 512], 86), ([4096], 344), ([704], 172)])
def test_gemm_lut_tma_sweep(fmt, dtype, t, shape):
    """TMA-fed table GEMV (pq_gemv_lut2.cu): ragged partitions and row blocks that are not a
    multiple of the 64-row box, K that ends inside a 64-position tile or a 16-byte chunk, code
    pitches that are not a multiple of 16 bytes (TMA super-rows of 2, 4, 8 rows - the TP = 4 / 8
    shards of down_proj), every
    cluster size the planner picks (1..8 tiles and more), several tiles per CTA, masked codes
    (nbits < 8), 1..8 tokens on grid.y; reproducible bit for bit."""
    b, d, m = fmt
    parts, groups = shape
    k = groups * d
    n = sum(parts)
    if not _lib.load().pq_b200_algo_supported(_lib.ALGO_GEMV_LUT_TMA, t, n, k, d, b, m):
        pytest.skip("shape outside the TMA kernel (odd pitch with too few rows)")
    c = O.make_case(parts, k, b, d, m, t=t, dtype=dtype, seed=7 + t, bias=(t % 2 == 0))
    g = dev(c)
    y = P.ops.pq_gemm(g["x"], g["codes"], g["codebooks"], g["scales"], parts, g["bias"], algo=_lib.ALGO_GEMV_LUT_TMA)
    check_output(y, c, "lut_tma")
    y2 = P.ops.pq_gemm(g["x"], g["codes"], g["codebooks"], g["scales"], parts, g["bias"], algo=_lib.ALGO_GEMV_LUT_TMA)
    assert torch.equal(y, y2)


MMA_SWEEP = [(8, 8, 1), (8, 8, 2), (4, 8, 1), (6, 8, 2)]


@pytest.mark.parametrize("fmt", MMA_SWEEP)
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("t", [1, 2, 5, 8, 9, 16, 23, 32, 40, 64])
@pytest.mark.parametrize("shape", [([128, 64, 72], 512), ([37], 8 * 48), ([1024, 48], 2048), ([200], 8 * 80)])
def test_gemm_mma_sweep(fmt, dtype, t, shape):
    """Register-fragment mma.sync kernel (2 <= T <= 64): ragged partitions, K that is not a
    multiple of a 64-byte superblock, every token-block count, bias on even T."""
    b, d, m = fmt
    parts, k = shape
    c = O.make_case(parts, k, b, d, m, t=t, dtype=dtype, seed=3 + t, bias=(t % 2 == 0))
    g = dev(c)
    assert _lib.load().pq_b200_algo_supported(_lib.ALGO_GEMM_MMA, t, sum(parts), k, d, b, m)
    y = P.ops.pq_gemm(g["x"], g["codes"], g["codebooks"], g["scales"], parts, g["bias"], algo=_lib.ALGO_GEMM_MMA)
    check_output(y, c, "mma")
    y2 = P.ops.pq_gemm(g["x"], g["codes"], g["codebooks"], g["scales"], parts, g["bias"], algo=_lib.ALGO_GEMM_MMA)
    assert torch.equal(y, y2)                    # fixed summation order


TC_SWEEP = [(8, 8, 1), (8, 8, 2), (16, 8, 1), (4, 4, 1), (8, 16, 1), (12, 8, 1), (8, 8, 3), (12, 4, 2)]


@pytest.mark.parametrize("fmt", TC_SWEEP)
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("t", [9, 16, 40, 130, 257])
def test_gemm_tc_sweep(fmt, dtype, t):
    """tcgen05 GEMM (prefill / batched decode): ragged partitions (tiles clipped at partition ends),
    token counts that are not multiples of the MMA N, bias on odd cases."""
    b, d, m = fmt
    parts, k = [128, 200, 72], 256
    c = O.make_case(parts, k, b, d, m, t=t, dtype=dtype, seed=11 + t, bias=(t % 2 == 1))
    g = dev(c)
    assert _lib.load().pq_b200_algo_supported(_lib.ALGO_GEMM_TC, t, sum(parts), k, d, b, m)
    y = P.ops.pq_gemm(g["x"], g["codes"], g["codebooks"], g["scales"], parts, g["bias"], algo=_lib.ALGO_GEMM_TC)
    check_output(y, c, "tc")
    # the A operand is the bit-exact dequantised weight and accumulation is fp32 in a fixed order:
    # the result is reproducible
    y2 = P.ops.pq_gemm(g["x"], g["codes"], g["codebooks"], g["scales"], parts, g["bias"], algo=_lib.ALGO_GEMM_TC)
    assert torch.equal(y, y2)
    check_output(P.ops.pq_gemm(g["x"], g["codes"], g["codebooks"], g["scales"], parts, g["bias"]), c, "auto")


def test_gemm_tc_prefill_full_size():
    """BASELINE config 3 shape class at reduced N (Llama-3-8B K, 2048 tokens): tolerance against a
    dense fp64 GEMM over pq_dequant's own (bit-exact) weights."""
    parts, k, t = [1024, 1024], 4096, 2048
    c = O.make_case(parts, k, 8, 8, 1, t=t, dtype=torch.bfloat16, seed=3)
    g = dev(c)
    y = P.ops.pq_gemm(g["x"], g["codes"], g["codebooks"], g["scales"], parts, None)
    w = P.ops.pq_dequant(g["codes"], g["codebooks"], parts).double() * g["scales"].double().view(-1, 1)
    ref = g["x"].double() @ w.t()
    rms = float((y.double() - ref).pow(2).mean().sqrt() / ref.pow(2).mean().sqrt())
    assert rms < 1.6e-2 / 4, rms


@pytest.mark.parametrize("shape", [
    # (parts, K) incl. rows whose code bytes are not a multiple of 16 (scalar fallback) and Llama TP=8 down
    ([37], 8 * 43), ([4096], 1376), ([48], 8 * 2), ([8, 8], 8 * 1), ([300, 4], 8 * 100)])
def test_gemm_ragged_shapes(shape):
    parts, k = shape
    for (b, d, m) in ((8, 8, 1), (8, 8, 2), (16, 8, 1)):
        c = O.make_case(parts, k, b, d, m, t=3, dtype=torch.float16, seed=9, bias=True)
        g = dev(c)
        y = P.ops.pq_gemm(g["x"], g["codes"], g["codebooks"], g["scales"], parts, g["bias"])
        check_output(y, c)


def test_leading_dims_and_strided_input():
    c = O.make_case([64, 64], 256, 8, 8, 1, t=6, dtype=torch.float16, seed=2)
    g = dev(c)
    y = P.ops.pq_gemm(g["x"].view(2, 3, 256), g["codes"], g["codebooks"], g["scales"], [64, 64], None)
    assert y.shape == (2, 3, 128)
    check_output(y.view(6, 128), c)
    wide = torch.zeros(6, 512, dtype=torch.float16, device="cuda")
    wide[:, :256] = g["x"]
    y2 = P.ops.pq_gemm(wide[:, :256], g["codes"], g["codebooks"], g["scales"], [64, 64], None)
    assert torch.equal(y2, y.view(6, 128))
    assert P.ops.pq_gemm(g["x"][:0], g["codes"], g["codebooks"], g["scales"], [64, 64], None).shape == (0, 128)


def test_errors_are_runtime_errors():
    c = dev(O.make_case([64], 64, 8, 8, 1, t=1, dtype=torch.float16))
    with pytest.raises(RuntimeError):      # three codebooks for two partitions: gemm_kernels.cu:539
        P.ops.pq_gemm(c["x"], c["codes"], c["codebooks"].repeat(3, 1, 1, 1), c["scales"], [32, 32], None)
    with pytest.raises(RuntimeError):
        P.ops.pq_gemm(c["x"][:, :32], c["codes"], c["codebooks"], c["scales"], [64], None)
    with pytest.raises(RuntimeError):
        P.ops.pq_gemm(c["x"].float(), c["codes"], c["codebooks"], c["scales"], [64], None)
    with pytest.raises(RuntimeError):
        P.ops.pq_dequant(c["codes"].to(torch.int16), c["codebooks"], [64])


def test_opcheck_like_reference():
    """tests/kernels/test_aqlm.py:9-39 - schema / faketensor / dispatch checks, reference shapes."""
    codes = torch.randint(-32768, 32767, (2048, 512, 1), device="cuda", dtype=torch.int16)
    codebooks = torch.rand((2, 65536, 1, 8), device="cuda", dtype=torch.float16)
    torch.library.opcheck(torch.ops.pq.dequant.default, (codes, codebooks, [1024, 1024]),
                          test_utils=("test_schema", "test_faketensor"))
    x = torch.rand((4, 4096), device="cuda", dtype=torch.float16)
    scales = torch.rand((2048, 1, 1, 1), device="cuda", dtype=torch.float16)
    torch.library.opcheck(torch.ops.pq.gemm.default, (x, codes, codebooks, scales, [1024, 1024], None),
                          test_utils=("test_schema", "test_faketensor"))


def test_cuda_graph_capture():
    c = dev(O.make_case([256, 256], 512, 8, 8, 2, t=2, dtype=torch.float16, seed=4))
    args = (c["x"], c["codes"], c["codebooks"], c["scales"], [256, 256], None)
    eager = P.ops.pq_gemm(*args)
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        P.ops.pq_gemm(*args)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            out = P.ops.pq_gemm(*args)
    g.replay()
    torch.cuda.synchronize()
    assert torch.equal(out, eager)


# ---- full-size, oracle-free properties (BASELINE.json sizes) -------------------------------
FULL = [("qkv", [4096] * 3, 4096), ("gate_up", [11008] * 2, 4096), ("down", [4096], 11008)]


@pytest.mark.parametrize("layer", FULL, ids=[f[0] for f in FULL])
@pytest.mark.parametrize("fmt", [(8, 8, 1), (8, 8, 2), (16, 8, 1)])
def test_full_size_properties(layer, fmt):
    _, parts, k = layer
    b, d, m = fmt
    c = O.make_case(parts, k, b, d, m, t=4, dtype=torch.float16, seed=1)
    g = dev(c)
    n = sum(parts)
    w = P.ops.pq_dequant(g["codes"], g["codebooks"], parts)
    # (1) dequant == plain gather of the codebooks, computed independently in torch on the GPU
    idx = (g["codes"].long() % (1 << b))
    acc = torch.zeros(n, k // d, d, dtype=torch.float32, device="cuda")
    off = 0
    for p, n_p in enumerate(parts):
        for j in range(m):
            acc[off:off + n_p] += g["codebooks"][p * m + j, :, 0, :].float()[idx[off:off + n_p, :, j]]
        off += n_p
    assert torch.equal(w, acc.to(torch.float16).view(n, k))
    # (2) a sample of rows against the CPU oracle, bit for bit
    rows = torch.arange(0, n, max(1, n // 64))
    sub = O.dequantize_partitions(c["codes"][rows], c["codebooks"][:m], [len(rows)], None) if len(parts) == 1 else None
    if sub is not None:
        assert G.same_bits(w[rows.cuda()], sub)
    # (3) the layer equals a dense fp32 GEMM over its own dequantised weight
    ws = P.ops.pq_dequant(g["codes"], g["codebooks"], parts, g["scales"])
    y = P.ops.pq_gemm(g["x"], g["codes"], g["codebooks"], g["scales"], parts, None)
    ref = (g["x"].double() @ (w.double() * g["scales"].double().view(n, 1)).t())
    rms = float(((y.double() - ref).pow(2).mean().sqrt() / ref.pow(2).mean().sqrt()))
    assert rms < 2e-3, rms
    for algo in (_lib.ALGO_GEMV_DIRECT, _lib.ALGO_GEMV_LUT, _lib.ALGO_GEMM_MMA, _lib.ALGO_GEMV_LUT_TMA):
        if algo != _lib.ALGO_GEMV_DIRECT and b > 8:
            continue
        ya = P.ops.pq_gemm(g["x"], g["codes"], g["codebooks"], g["scales"], parts, None, algo=algo)
        rms = float(((ya.double() - ref).pow(2).mean().sqrt() / ref.pow(2).mean().sqrt()))
        assert rms < 2e-3, (algo, rms)
    assert torch.isfinite(ws).all()
    # (4) linearity in x: f(2x) == 2 f(x) exactly (power-of-two scaling commutes with rounding)
    y2 = P.ops.pq_gemm(g["x"] * 2, g["codes"], g["codebooks"], g["scales"], parts, None)
    normal = y.abs() > 1e-3                      # fp16 subnormal outputs round differently when doubled
    assert torch.equal(y2[normal], (y * 2)[normal])
