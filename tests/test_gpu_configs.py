"""BASELINE.json's configurations at their OWN sizes, against the oracle.

The CPU oracle evaluates the fp64 contraction on a sample of rows of every output partition
(first, last and ~64 evenly spaced rows: the codebooks, x and every token are the full-size ones),
so a 57344 x 8192 layer costs milliseconds of host time; the CUDA result is the full layer.
  config 1   single 4096 x 4096 linear, 8 bit x d8 x M1, one token
  config 2   the four Llama-2-7B projections at T = 16, 17, 32, 64 (where PQ_ALGO_AUTO switches kernels)
  config 3   Llama-3-8B MLP shapes, 2048 tokens
  config 4   the Llama-2-70B TP = 2 / 4 / 8 shard shapes of SURVEY 8a A13, on one GPU
  strided x  the table GEMV fed a row-strided x at full size (what the decoder stack hands it)
"""
import pytest
import torch

import vllm_pq_b200 as P
from oracle import pq_oracle as O
from tests.test_gpu_parity import check_output, dev
from vllm_pq_b200 import _lib

pytestmark = pytest.mark.gpu

FORMATS = [(8, 8, 1), (8, 8, 2), (16, 8, 1)]


def row_sample(c, per_part=64):
    """(sub-case over a sample of rows of every partition, their indices in the full layer)."""
    sel, parts, off = [], [], 0
    for n_p in c["partition_sizes"]:
        r = torch.arange(off, off + n_p, max(1, n_p // per_part))
        r = torch.unique(torch.cat([r, torch.tensor([off, off + n_p - 1])]))
        sel.append(r)
        parts.append(len(r))
        off += n_p
    sel = torch.cat(sel)
    sub = dict(c)
    sub["codes"], sub["scales"] = c["codes"][sel].contiguous(), c["scales"][sel].contiguous()
    sub["bias"] = None if c["bias"] is None else c["bias"][sel].contiguous()
    sub["partition_sizes"] = parts
    return sub, sel


def check_sampled(y, c, name=""):
    sub, sel = row_sample(c)
    assert torch.isfinite(y).all(), name
    return check_output(y.reshape(-1, y.shape[-1])[:, sel.to(y.device)], sub, name)


def gemm(g, parts, algo=_lib.ALGO_AUTO, x=None):
    return P.ops.pq_gemm(g["x"] if x is None else x, g["codes"], g["codebooks"], g["scales"], parts, g["bias"], algo=algo)


# ---- config 1 ------------------------------------------------------------------------------
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
def test_config1_single_linear(dtype):
    c = O.make_case([4096], 4096, 8, 8, 1, t=1, dtype=dtype, seed=11)
    g = dev(c)
    lib = _lib.load()
    assert lib.pq_b200_pick_algo(1, 4096, 4096, 8, 8, 1) == _lib.ALGO_GEMV_LUT_TMA
    y = gemm(g, [4096])
    check_output(y, c, "auto")                               # every row, not a sample
    for algo in (_lib.ALGO_GEMV_DIRECT, _lib.ALGO_GEMV_LUT, _lib.ALGO_GEMM_MMA):
        check_output(gemm(g, [4096], algo), c, _lib.ALGO_NAMES[algo])
    w = P.ops.pq_dequant(g["codes"], g["codebooks"], [4096], g["scales"])
    w_ref = O.dequantize_partitions(c["codes"], c["codebooks"], [4096], c["scales"])
    assert torch.equal(w.cpu().view(torch.int16), w_ref.view(torch.int16))


# ---- strided x into the table GEMV at full size ------------------------------------------------
@pytest.mark.parametrize("layer", [("down7b", [4096], 11008, 22016), ("o7b", [4096], 4096, 12288)])
@pytest.mark.parametrize("t", [1, 4])
def test_strided_x_full_size(layer, t):
    """The decoder stack hands o_proj / down_proj a slice of the wider qkv / gate_up output:
    x_stride 12288 / 22016 elements with K = 4096 / 11008."""
    _, parts, k, stride = layer
    c = O.make_case(parts, k, 8, 8, 1, t=t, dtype=torch.float16, seed=12)
    g = dev(c)
    wide = torch.randn(t, stride, dtype=torch.float16, device="cuda")
    wide[:, :k] = g["x"]
    xs = wide[:, :k]
    assert xs.stride(0) == stride and not xs.is_contiguous() or t == 1
    for algo in (_lib.ALGO_AUTO, _lib.ALGO_GEMV_LUT_TMA):
        y = gemm(g, parts, algo, x=xs)
        assert torch.equal(y, gemm(g, parts, algo)), "strided x changed the result"
        check_output(y, c, f"strided/{_lib.ALGO_NAMES[algo]}")
    # offset view (column slice that does not start at 0: the 'up' half of gate_up), 16-byte aligned
    wide2 = torch.randn(t, 2 * k, dtype=torch.float16, device="cuda")
    wide2[:, k:] = g["x"]
    assert torch.equal(gemm(g, parts, x=wide2[:, k:]), gemm(g, parts))


# ---- config 2 at the token counts where AUTO changes kernel ------------------------------------
PROJ7B = [("qkv", [4096] * 3, 4096), ("o", [4096], 4096), ("gate_up", [11008] * 2, 4096), ("down", [4096], 11008)]


@pytest.mark.parametrize("layer", PROJ7B, ids=[p[0] for p in PROJ7B])
@pytest.mark.parametrize("fmt", FORMATS, ids=["8x8x1", "8x8x2", "16x8x1"])
def test_config2_token_counts(layer, fmt):
    _, parts, k = layer
    b, d, m = fmt
    c = O.make_case(parts, k, b, d, m, t=64, dtype=torch.float16, seed=13)
    g = dev(c)
    y64 = None
    for t in (16, 17, 32, 64):
        ct = dict(c)
        ct["x"] = c["x"][:t]
        y = gemm(g, parts, x=g["x"][:t])
        check_sampled(y, ct, f"t={t}")
        if t == 64:
            y64 = y
    # a token's result does not depend on how many tokens share the launch beyond rounding of the
    # kernel family: compare one-token GEMV against row 0 of the 64-token result within tolerance
    y1 = gemm(g, parts, x=g["x"][:1])
    rel = float((y1.double() - y64[:1].double()).pow(2).mean().sqrt() / y64[:1].double().pow(2).mean().sqrt())
    assert rel < 4e-3, rel


# ---- config 3 at its real N ----------------------------------------------------------------------
MLP8B = [("gate_up8b", [14336] * 2, 4096), ("down8b", [4096], 14336)]


@pytest.mark.parametrize("layer", MLP8B, ids=[p[0] for p in MLP8B])
@pytest.mark.parametrize("fmt", FORMATS, ids=["8x8x1", "8x8x2", "16x8x1"])
def test_config3_prefill_real_size(layer, fmt):
    _, parts, k = layer
    b, d, m = fmt
    c = O.make_case(parts, k, b, d, m, t=2048, dtype=torch.bfloat16, seed=14)
    g = dev(c)
    lib = _lib.load()
    assert lib.pq_b200_pick_algo(2048, sum(parts), k, d, b, m) == _lib.ALGO_GEMM_TC
    y = gemm(g, parts)
    check_sampled(y, c, "prefill")
    assert torch.equal(y, gemm(g, parts)), "prefill GEMM is not reproducible"


# ---- config 4: Llama-2-70B shard shapes on one GPU -------------------------------
def shards_70b(tp):
    q, kv = 8192 // tp, max(1, 8 // tp) * 128
    return [("qkv", [q, kv, kv], 8192), ("o", [8192], 8192 // tp),
            ("gate_up", [28672 // tp] * 2, 8192), ("down", [8192], 28672 // tp)]


@pytest.mark.parametrize("tp", [2, 4, 8])
@pytest.mark.parametrize("fmt", FORMATS, ids=["8x8x1", "8x8x2", "16x8x1"])
def test_config4_70b_shard_shapes(tp, fmt):
    b, d, m = fmt
    for name, parts, k in shards_70b(tp):
        c = O.make_case(parts, k, b, d, m, t=16, dtype=torch.float16, seed=15 + tp)
        g = dev(c)
        for t in (1, 16):
            ct = dict(c)
            ct["x"] = c["x"][:t]
            check_sampled(gemm(g, parts, x=g["x"][:t]), ct, f"tp{tp}/{name}/t={t}")


# ---- unaligned code pitch (Llama-2-7B down_proj shards at TP = 4 / 8: 344 / 172 code bytes per row) at
# every token regime: the table GEMV's cp.async loader (1..8), the tcgen05 kernel's generic code path (9+)
@pytest.mark.parametrize("k", [2752, 1376])
@pytest.mark.parametrize("fmt", [(8, 8, 1), (8, 8, 2)], ids=["8x8x1", "8x8x2"])
def test_unaligned_pitch_all_token_counts(k, fmt):
    b, d, m = fmt
    c = O.make_case([4096], k, b, d, m, t=300, dtype=torch.float16, seed=17)
    g = dev(c)
    for t in (1, 2, 8, 9, 16, 64, 300):
        ct = dict(c)
        ct["x"] = c["x"][:t]
        check_sampled(gemm(g, [4096], x=g["x"][:t]), ct, f"k={k} t={t}")


@pytest.mark.parametrize("layer", MLP8B + [("qkv_parts", [1024, 128, 128], 4096)], ids=["gate_up8b", "down8b", "qkv_parts"])
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
def test_prefill_16bit_dense_weight_path(layer, dtype, monkeypatch):
    """16-bit codes at >= 1024 tokens: the weight is dequantised once into the workspace and TMA-loaded as the
    tcgen05 A operand.  Same operand bits, same accumulation order as the in-kernel dequant -> the SAME
    output bits; and the ticket area at the head of the shared workspace stays zero (a K-group table
    GEMV launched afterwards on the same stream still works)."""
    _, parts, k = layer
    c = O.make_case(parts, k, 16, 8, 1, t=1024, dtype=dtype, seed=19)
    g = dev(c)
    y = gemm(g, parts, _lib.ALGO_GEMM_TC)
    check_sampled(y, c, "dense-A prefill")
    # a table GEMV with K groups through the same workspace afterwards
    c8 = O.make_case([4096], 11008, 8, 8, 1, t=1, dtype=dtype, seed=20)
    g8 = dev(c8)
    check_output(gemm(g8, [4096], _lib.ALGO_GEMV_LUT_TMA), c8, "table GEMV after the dense prefill")
    # 1023 tokens take the in-kernel dequant: row 0..1022 must carry the same bits
    y2 = gemm(g, parts, _lib.ALGO_GEMM_TC, x=g["x"][:1023])
    assert torch.equal(y2, y[:1023])


# ---- Llama-2-7B TP = 2 / 4 / 8 shard shapes through the table GEMV (what the SCALE run launches) -------------
def shards_7b(tp):
    return [("qkv", [4096 // tp] * 3, 4096), ("o", [4096], 4096 // tp),
            ("gate_up", [11008 // tp] * 2, 4096), ("down", [4096], 11008 // tp)]


@pytest.mark.parametrize("tp", [2, 4, 8])
@pytest.mark.parametrize("fmt", [(8, 8, 1), (8, 8, 2)], ids=["8x8x1", "8x8x2"])
def test_7b_tp_shard_shapes_table_gemv(tp, fmt):
    """Every plan the planner picks for the TP shards (16-CTA clusters of the 8-warp shape with two row
    iterations for down_proj at TP = 2, clusters wider than the tile count at TP = 4 / 8, the cp.async
    loader for unaligned pitches): against the oracle, reproducible bit for bit, 1 / 2 / 8 tokens."""
    b, d, m = fmt
    lib = _lib.load()
    for name, parts, k in shards_7b(tp):
        c = O.make_case(parts, k, b, d, m, t=8, dtype=torch.float16, seed=23 + tp)
        g = dev(c)
        for t in (1, 2, 8):
            if not lib.pq_b200_algo_supported(_lib.ALGO_GEMV_LUT_TMA, t, sum(parts), k, d, b, m):
                continue
            ct = dict(c)
            ct["x"] = c["x"][:t]
            y = gemm(g, parts, _lib.ALGO_GEMV_LUT_TMA, x=g["x"][:t])
            check_output(y, ct, f"tp{tp}/{name}/t={t}")
            assert torch.equal(y, gemm(g, parts, _lib.ALGO_GEMV_LUT_TMA, x=g["x"][:t])), (tp, name, t)
