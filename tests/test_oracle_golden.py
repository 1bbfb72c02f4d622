"""The oracle against the reference: committed golden vectors (always) and the reference's
own aqlm.py loaded from /root/reference (only where that tree exists)."""
import pytest
import torch

from oracle import pq_oracle as O
from oracle import ref_loader
from tests import golden_util as G


@pytest.mark.parametrize("name", G.names())
def test_oracle_matches_golden(name):
    c = G.load(name)
    parts = c["partition_sizes"]
    wu = O.dequantize_partitions(c["codes"], c["codebooks"], parts, None)
    ws = O.dequantize_partitions(c["codes"], c["codebooks"], parts, c["scales"])
    assert G.same_bits(wu, c["w_unscaled"])            # P1, bit-exact
    assert G.same_bits(ws, c["w_scaled"])              # P2, bit-exact
    y = O.generic_dequantize_gemm(c["x"], c["codes"], c["codebooks"], c["scales"], parts, c["bias"])
    assert G.same_bits(y, c["y"])                      # same torch ops on the same weights
    # the literal embedding_bag formulation (used for the timed CPU baseline) agrees too (fp16)
    if c["dtype"] == torch.float16:
        y2 = O.generic_dequantize_gemm(c["x"], c["codes"], c["codebooks"], c["scales"], parts,
                                       c["bias"], literal=True)
        assert G.same_bits(y2, c["y"])


def test_golden_covers_corner_codes():
    c = G.load("c1x8_d8")
    assert c["codes"].view(-1)[:4].tolist() == [-1, -128, 127, 0]
    idx = O.unpack_int_data(c["codes"], 8).view(-1)[:4].tolist()
    assert idx == [255, 128, 127, 0]                   # aqlm.py:32-34


def test_unpack_is_mod_2b():
    import numpy as np
    a = np.array([-8, -1, 0, 7], dtype=np.int8)
    assert O.unpack_codes_np(a, 4).tolist() == [8, 15, 0, 7]
    a = np.array([-32768, -1, 32767], dtype=np.int16)
    assert O.unpack_codes_np(a, 16).tolist() == [32768, 65535, 32767]
    a = np.array([-2048, -1, 2047], dtype=np.int16)
    assert O.unpack_codes_np(a, 12).tolist() == [2048, 4095, 2047]


def test_int_dtype_table():
    assert O.get_int_dtype(4) == torch.int8 and O.get_int_dtype(8) == torch.int8
    assert O.get_int_dtype(9) == torch.int16 and O.get_int_dtype(16) == torch.int16
    assert O.get_int_dtype(17) == torch.int32
    with pytest.raises(ValueError):
        O.get_int_dtype(65)


@pytest.mark.skipif(not ref_loader.available(), reason="/root/reference not present (GPU box)")
@pytest.mark.parametrize("fmt", [(8, 8, 1), (8, 8, 2), (16, 8, 1), (4, 4, 1), (12, 16, 1), (8, 8, 4),
                                 (8, 8, 8), (4, 8, 2), (8, 16, 1), (12, 4, 1)])
def test_oracle_matches_live_reference(fmt):
    ref = ref_loader.load_reference_aqlm()
    b, d, m = fmt
    parts = [40, 24, 24]
    c = O.make_case(parts, 16 * d, b, d, m, t=3, dtype=torch.float16, seed=7, bias=True,
                    positive_scales=False)
    off = 0
    for p, n_p in enumerate(parts):
        cu = ref.unpack_int_data(c["codes"][off:off + n_p], b)
        for s in (None, c["scales"][off:off + n_p]):
            wr = ref.dequantize_weight(cu, c["codebooks"][p * m:(p + 1) * m], s)
            wo = O.dequantize_weight(cu, c["codebooks"][p * m:(p + 1) * m], s)
            assert G.same_bits(wr, wo)
        off += n_p
    yr = ref.generic_dequantize_gemm(c["x"], c["codes"], c["codebooks"], c["scales"], parts, c["bias"])
    yo = O.generic_dequantize_gemm(c["x"], c["codes"], c["codebooks"], c["scales"], parts, c["bias"])
    assert G.same_bits(yr, yo)
    assert ref.get_int_dtype(b) == O.get_int_dtype(b)


@pytest.mark.skipif(not ref_loader.available(), reason="/root/reference not present (GPU box)")
def test_reference_bf16_single_codebook_matches():
    """bf16 is an extension (reference: fp16 only, aqlm.py:238-239). For M=1 the reference's torch
    functions still run and agree with the oracle; for M>=2 CPU embedding_bag does not round to
    nearest even in bf16, so the oracle's fp32-sum + RNE is the definition."""
    ref = ref_loader.load_reference_aqlm()
    c = O.make_case([64], 128, 8, 8, 1, t=2, dtype=torch.bfloat16, seed=3)
    cu = ref.unpack_int_data(c["codes"], 8)
    assert G.same_bits(ref.dequantize_weight(cu, c["codebooks"], c["scales"]),
                       O.dequantize_weight(cu, c["codebooks"], c["scales"]))


# ---- the restatements behind the fused epilogue against the reference's layers ----
import os

import numpy as np

LAYER_DIR = os.path.join(G.GOLDEN_DIR, "layers")


def _layer_fixture(name):
    z = np.load(os.path.join(LAYER_DIR, name + ".npz"))
    dtype = torch.bfloat16 if bool(z["is_bf16"]) else torch.float16
    out = {k: torch.from_numpy(z[k].copy()).view(dtype) for k in z.files if k not in ("eps", "is_bf16")}
    if "eps" in z.files:
        out["eps"] = float(z["eps"])
    return out


def _ulps(a, b):
    """Distance in units of the last place between two tensors of one 16-bit float type."""
    def key(t):
        i = t.view(torch.int16).to(torch.int32)
        return torch.where(i < 0, -(i & 0x7FFF), i)
    return (key(a) - key(b)).abs()


@pytest.mark.parametrize("tag", ["f16", "bf16"])
def test_rms_norm_restatement_matches_reference_layer(tag):
    """RMSNorm.forward_native (layernorm.py:38-74), committed outputs.  The new residual is bit-exact.
    The normalised output is the same two roundings; the variance is summed in fp64 here and in fp32
    there, so rstd may differ in its last bit: that can move the first rounding by one ulp, which the
    weight (|w| up to ~1.6 here) stretches to at most two in the second - on well under 1 % of the elements."""
    f = _layer_fixture("norm_" + tag)
    out, z = O.fused_add_rms_norm(f["x"], None, f["weight"], f["eps"])
    assert z is f["x"]
    d = _ulps(out, f["out"])
    assert int(d.max()) <= 2 and float((d > 0).float().mean()) < 0.01
    out2, z2 = O.fused_add_rms_norm(f["x"], f["residual"], f["weight"], f["eps"])
    assert G.same_bits(z2, f["z"])
    # the CUDA kernels (the reference's and ours) normalise the ROUNDED sum: forward_native of z
    d = _ulps(out2, f["out_z"])
    assert int(d.max()) <= 2 and float((d > 0).float().mean()) < 0.01
    # forward_native(x, residual) keeps the fp32 sum: further away, and that is the reference's own
    # CUDA-vs-native difference, not the restatement's (documented in make_golden_layers.py)
    far = _ulps(out2, f["out_res"])
    assert int(far.max()) <= 3 and float((far > 0).float().mean()) > 0.05


@pytest.mark.parametrize("tag", ["f16", "bf16"])
def test_silu_and_mul_restatement_matches_reference_layer(tag):
    f = _layer_fixture("silu_" + tag)
    assert G.same_bits(O.silu_and_mul(f["x"]), f["out"])


@pytest.mark.skipif(not ref_loader.available(), reason="/root/reference not present (GPU box)")
def test_layer_restatements_match_live_reference():
    RMSNorm, SiluAndMul = ref_loader.load_reference_layers()
    g = torch.Generator().manual_seed(9)
    for dtype in (torch.float16, torch.bfloat16):
        x = torch.randn(5, 1024, generator=g).to(dtype)
        r = torch.randn(5, 1024, generator=g).to(dtype)
        norm = RMSNorm(1024, 1e-6)
        norm.weight.data = (1.0 + 0.1 * torch.randn(1024, generator=g)).to(dtype)
        with torch.no_grad():
            _, z_ref = norm.forward_native(x, r)
            out_ref = norm.forward_native(z_ref)
        out, z = O.fused_add_rms_norm(x, r, norm.weight.data, 1e-6)
        assert G.same_bits(z, z_ref) and int(_ulps(out, out_ref).max()) <= 2
        gu = torch.randn(3, 512, generator=g).to(dtype)
        assert G.same_bits(O.silu_and_mul(gu), SiluAndMul().forward_native(gu))


# ---- random formats and shapes: the restatement against the live reference ----------------------------
from hypothesis import given, settings
from hypothesis import strategies as st


@pytest.mark.skipif(not ref_loader.available(), reason="/root/reference not present (GPU box)")
@settings(max_examples=40, deadline=None, derandomize=True)
@given(b=st.integers(1, 16), d=st.sampled_from([1, 2, 4, 8, 16, 24]), m=st.integers(1, 6),
       parts=st.lists(st.integers(1, 24), min_size=1, max_size=4), groups=st.integers(1, 12),
       t=st.integers(1, 5), bias=st.booleans(), seed=st.integers(0, 1000))
def test_oracle_matches_live_reference_on_random_formats(b, d, m, parts, groups, t, bias, seed):
    """Any (nbits, d, M), any partition list, ragged K: dequantised weights (unscaled and scaled) and the
    generic GEMM carry the bits the reference's functions return (fp16: its only dtype, aqlm.py:238-239)."""
    ref = ref_loader.load_reference_aqlm()
    c = O.make_case(parts, groups * d, b, d, m, t=t, dtype=torch.float16, seed=seed, bias=bias, positive_scales=False)
    assert ref.get_int_dtype(b) == O.get_int_dtype(b) == c["codes"].dtype
    off = 0
    for p, n_p in enumerate(parts):
        cu = ref.unpack_int_data(c["codes"][off:off + n_p], b)
        assert torch.equal(cu, O.unpack_int_data(c["codes"][off:off + n_p], b))
        for s in (None, c["scales"][off:off + n_p]):
            assert G.same_bits(ref.dequantize_weight(cu, c["codebooks"][p * m:(p + 1) * m], s),
                               O.dequantize_weight(cu, c["codebooks"][p * m:(p + 1) * m], s))
        off += n_p
    assert G.same_bits(ref.generic_dequantize_gemm(c["x"], c["codes"], c["codebooks"], c["scales"], parts, c["bias"]),
                       O.generic_dequantize_gemm(c["x"], c["codes"], c["codebooks"], c["scales"], parts, c["bias"]))
