"""Host-side logic, no GPU: the C-ABI library loads and exports what include/pq_b200.h declares,
the plugin API mirrors the reference's, and the product never routes through the oracle."""
import ctypes
import os
import re

import pytest
import torch

import vllm_pq_b200 as P
from oracle import ref_loader
from vllm_pq_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    text = open(os.path.join(ROOT, "include", "pq_b200.h")).read()
    return sorted(set(re.findall(r"PQ_API\s+[\w\s\*]+?\b(pq_b200_\w+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    lib = ctypes.CDLL(_lib.LIB_PATH)
    syms = header_symbols()
    assert len(syms) >= 14
    for s in syms:
        assert hasattr(lib, s), s
    assert sorted(_lib.SIGNATURES) == syms          # python binding covers the whole header
    assert _lib.load().pq_b200_abi_version() == _lib.ABI_VERSION


def test_argument_errors_without_gpu():
    lib = _lib.load()
    parts, n_parts = _lib.parts_array([16])
    # bad dtype / bad format / bad shape are rejected before any launch
    assert lib.pq_b200_dequant(1, 16, None, 16, 16, 64, 8, 8, 1, parts, n_parts, 7, None) == -1
    assert lib.pq_b200_dequant(1, 16, None, 16, 16, 64, 8, 17, 1, parts, n_parts, 0, None) == -2
    assert lib.pq_b200_dequant(1, 16, None, 16, 16, 63, 8, 8, 1, parts, n_parts, 0, None) == -3
    assert lib.pq_b200_dequant(1, 16, None, 16, 17, 64, 8, 8, 1, parts, n_parts, 0, None) == -3
    assert lib.pq_b200_dequant(None, 16, None, 16, 16, 64, 8, 8, 1, parts, n_parts, 0, None) == -6
    assert lib.pq_b200_dequant(1, 8, None, 16, 16, 64, 8, 8, 1, parts, n_parts, 0, None) == -4
    assert b"partition" in lib.pq_b200_strerror(-3)
    with pytest.raises(RuntimeError):
        _lib.check(-3, "x")


def test_no_cpu_path():
    c = P.PQConfig(8, 8, 1, 1)
    layer = P.ReplicatedLinear(64, 32, bias=False, params_dtype=torch.float16, quant_config=c)
    with pytest.raises((RuntimeError, NotImplementedError)):
        layer(torch.zeros(1, 64, dtype=torch.float16))
    src = ""
    for dirpath, _, files in os.walk(os.path.join(ROOT, "vllm_pq_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh")):
                src += open(os.path.join(dirpath, f)).read()
    assert not re.search(r"^\s*(from|import)\s+oracle", src, re.M), "product must not import oracle/"


def test_config_mirrors_reference_contract():
    c = P.PQConfig.from_config({"in_group_size": 8, "nbits_per_codebook": 16, "num_codebooks": 1,
                                "out_group_size": 1, "quant_method": "aqlm"})
    assert (c.in_group_size, c.nbits_per_codebook, c.num_codebooks, c.out_group_size) == (8, 16, 1, 1)
    assert c.pack_factor == 8 and c.get_name() == "pq"
    assert torch.half in c.get_supported_act_dtypes() and torch.bfloat16 in c.get_supported_act_dtypes()
    assert P.PQConfig.get_min_capability() == 100 and P.PQConfig.get_config_filenames() == []
    with pytest.raises(ValueError):
        P.PQConfig.from_config({"in_group_size": 8})
    with pytest.raises(AssertionError):
        P.PQConfig(8, 8, 1, 2)                       # aqlm.py:178-179
    assert c.get_quant_method(torch.nn.Linear(4, 4), "") is None      # aqlm.py:213-217
    assert P.get_quantization_config("pq") is P.PQConfig
    with pytest.raises(ValueError):                  # test_register_quantization_config.py:49-97
        P.register_quantization_config("pq")(P.PQConfig)
    with pytest.raises(ValueError):
        P.register_quantization_config("other")(object)


FORMATS = [(8, 8, 1), (8, 8, 2), (16, 8, 1), (4, 4, 1), (12, 16, 1)]


@pytest.mark.parametrize("fmt", FORMATS)
@pytest.mark.parametrize("parts", [[4096, 4096, 4096], [11008, 11008], [4096]])
def test_create_weights_layout(fmt, parts):
    b, d, m = fmt
    method = P.PQLinearMethod(P.PQConfig(d, b, m, 1))
    layer = torch.nn.Module()
    with torch.device("meta"):
        method.create_weights(layer, 4096, parts, 4096, sum(parts), torch.float16, weight_loader=print)
    n = sum(parts)
    assert layer.codes.shape == (n, 4096 // d, m) and layer.codes.dtype == P.get_int_dtype(b)
    assert layer.codebooks.shape == (m * len(parts), 2**b, 1, d) and layer.codebooks.dtype == torch.float16
    assert layer.scales.shape == (n, 1, 1, 1)
    assert (layer.codes.input_dim, layer.codes.output_dim, layer.codes.packed_dim, layer.codes.pack_factor) == (1, 0, 1, d)
    assert layer.codebooks.is_metadata and layer.codebooks.output_partition_sizes == parts
    assert (layer.scales.output_dim, layer.scales.packed_dim, layer.scales.pack_factor) == (0, 0, 1)
    assert all(p.weight_loader is print and not p.requires_grad for p in (layer.codes, layer.codebooks, layer.scales))

    if ref_loader.available():                       # same names / shapes / dtypes / attrs as aqlm.py:253-323
        ref = ref_loader.load_reference_aqlm()
        rl = torch.nn.Module()
        with torch.device("meta"):
            ref.AQLMLinearMethod(ref.AQLMConfig(d, b, m, 1)).create_weights(
                rl, 4096, parts, 4096, n, torch.float16, weight_loader=print)
        for name in ("codes", "codebooks", "scales"):
            a, r = getattr(layer, name), getattr(rl, name)
            assert a.shape == r.shape and a.dtype == r.dtype, name
            for attr in ("input_dim", "output_dim", "packed_dim", "pack_factor", "is_metadata",
                         "output_partition_sizes", "weight_loader"):
                assert getattr(a, attr, None) == getattr(r, attr, None), (name, attr)
        assert [n_ for n_, _ in layer.named_parameters()] == [n_ for n_, _ in rl.named_parameters()]


def test_create_weights_errors():
    method = P.PQLinearMethod(P.PQConfig(8, 8, 1, 1))
    with pytest.raises(ValueError):                  # aqlm.py:238-239 (extended to bf16)
        method.create_weights(torch.nn.Module(), 64, [8], 64, 8, torch.float32)
    with pytest.raises(ValueError):                  # aqlm.py:240-244
        method.create_weights(torch.nn.Module(), 60, [8], 60, 8, torch.float16)


def test_fake_ops_trace_shapes():
    """Meta tensors go through the fake impls (vllm/_custom_ops.py:347-371) - what torch.compile sees."""
    x = torch.empty(2, 3, 4096, dtype=torch.float16, device="meta")
    codes = torch.empty(12288, 512, 1, dtype=torch.int16, device="meta")
    cb = torch.empty(3, 65536, 1, 8, dtype=torch.float16, device="meta")
    sc = torch.empty(12288, 1, 1, 1, dtype=torch.float16, device="meta")
    y = torch.ops.pq.gemm(x, codes, cb, sc, [4096] * 3, None)
    assert y.shape == (2, 3, 12288) and y.dtype == torch.float16
    w = torch.ops.pq.dequant(codes, cb, [4096] * 3)
    assert w.shape == (12288, 4096) and w.dtype == torch.float16
    # schema check of the reference's opcheck (tests/kernels/test_aqlm.py:9-39), minus the CUDA run
    torch.library.opcheck(torch.ops.pq.dequant.default, (codes, cb, [4096] * 3), test_utils=("test_schema",))


def test_tp1_layers_build():
    cfg = P.PQConfig(8, 8, 2, 1)
    qkv = P.QKVParallelLinear(4096, 128, 32, 32, bias=False, params_dtype=torch.float16, quant_config=cfg)
    assert qkv.output_partition_sizes == [4096, 4096, 4096] and qkv.codes.shape == (12288, 512, 2)
    assert qkv.codebooks.shape == (6, 256, 1, 8)
    gu = P.MergedColumnParallelLinear(4096, [11008, 11008], bias=False, params_dtype=torch.float16, quant_config=cfg)
    assert gu.output_partition_sizes == [11008, 11008]
    down = P.RowParallelLinear(11008, 4096, bias=False, params_dtype=torch.float16, quant_config=cfg)
    assert down.codes.shape == (4096, 1376, 2) and down.input_size_per_partition == 11008
    # fused-shard loaders: codebooks land at shard_index * M (linear.py:619-624, 1038-1044)
    cbk = torch.randn(2, 256, 1, 8).half()
    qkv.weight_loader(qkv.codebooks, cbk, "k")
    assert torch.equal(qkv.codebooks.data[2:4], cbk)
    codes_v = torch.randint(-128, 127, (4096, 512, 2), dtype=torch.int8)
    qkv.weight_loader(qkv.codes, codes_v, "v")
    assert torch.equal(qkv.codes.data[8192:], codes_v)
    up = torch.randn(11008, 1, 1, 1).half()
    gu.weight_loader(gu.scales, up, 1)
    assert torch.equal(gu.scales.data[11008:], up)


def test_silu_mul_ops_fake_shapes_and_hook_checks():
    """the fused-activation ops trace like the plain ones (gate|up in, N out), and
    RowParallelLinear refuses activations it does not know or an input it would have to slice."""
    x2 = torch.empty(2, 3, 2 * 11008, dtype=torch.float16, device="meta")
    codes = torch.empty(4096, 1376, 1, dtype=torch.int8, device="meta")
    cb = torch.empty(1, 256, 1, 8, dtype=torch.float16, device="meta")
    sc = torch.empty(4096, 1, 1, 1, dtype=torch.float16, device="meta")
    y = torch.ops.pq.gemm_silu_mul(x2, codes, cb, sc, [4096], None, 0)
    assert y.shape == (2, 3, 4096) and y.dtype == torch.float16
    a = torch.ops.pq.silu_and_mul(x2)
    assert a.shape == (2, 3, 11008) and a.dtype == torch.float16
    torch.library.opcheck(torch.ops.pq.silu_and_mul.default, (x2,), test_utils=("test_schema",))

    cfg = P.PQConfig(8, 8, 1, 1)
    down = P.RowParallelLinear(256, 128, bias=False, params_dtype=torch.float16, quant_config=cfg)
    with pytest.raises(ValueError, match="unknown input activation"):
        down(torch.zeros(1, 512, dtype=torch.float16), input_act="gelu")
    sliced = P.RowParallelLinear(256, 128, bias=False, input_is_parallel=False, params_dtype=torch.float16,
                                 quant_config=cfg)
    with pytest.raises(ValueError, match="input_is_parallel"):
        sliced(torch.zeros(1, 512, dtype=torch.float16), input_act="silu_and_mul")
    # no CPU fallback: the fused call on CPU tensors fails loudly, like the plain one
    with pytest.raises((RuntimeError, NotImplementedError)):
        down(torch.zeros(1, 512, dtype=torch.float16), input_act="silu_and_mul")
    # the oracle's activation is the reference's forward_native (activation.py:73-76)
    from oracle import pq_oracle as O
    v = torch.tensor([[1.0, -2.0, 0.5, 3.0]], dtype=torch.float16)
    assert torch.equal(O.silu_and_mul(v), torch.nn.functional.silu(v[:, :2]) * v[:, 2:])


def test_round2_entry_points_reject_bad_arguments_without_gpu():
    """The new C entry points validate before any CUDA call (include/pq_b200.h): fused GEMV, norm,
    embedding, two-stage all-reduce; and the fused-support predicate follows its contract."""
    lib = _lib.load()
    parts, n_parts = _lib.parts_array([4096])
    sup = lib.pq_b200_gemm_fused_supported
    assert sup(1, 4096, 4096, 8, 8, 1) == 1 and sup(8, 4096, 4096, 8, 8, 2) == 1
    assert sup(9, 4096, 4096, 8, 8, 1) == 0            # more than 8 tokens: separate launches
    assert sup(1, 4096, 4096, 8, 4, 1) == 0            # 4-bit codes as stored (the nibble-packed copy is 8-bit)
    assert sup(1, 4096, 4096, 8, 16, 1) == 0           # 16-bit codes: no table GEMV
    assert sup(1, 4096, 4095, 8, 8, 1) == 0
    assert lib.pq_b200_gemm_fused_slots(1, 4096, 4096, 8, 8, 1, parts, n_parts, 0) > 0
    assert lib.pq_b200_gemm_fused_slots(1, 4096, 4096, 8, 16, 1, parts, n_parts, 0) == -7
    ep = _lib.Epilogue()
    gf = lib.pq_b200_gemm_fused
    args = lambda **kw: [kw.get("x", 16), 4096, 16, 16, 16, None, 16, 4096, kw.get("t", 1), 4096, 4096, 8,   # noqa: E731
                         kw.get("nbits", 8), 1, parts, n_parts, kw.get("dtype", 0), None, 0, None, ctypes.byref(ep)]
    assert gf(*args(dtype=5)) == -1
    assert gf(*args(x=None)) == -6
    assert gf(*args(nbits=16)) == -7
    assert gf(*args(t=9)) == -7
    norm = lib.pq_b200_fused_add_rms_norm
    assert norm(16, 64, None, 0, 16, 1e-5, 1, 64, 3, None) == -1
    assert norm(None, 64, None, 0, 16, 1e-5, 1, 64, 0, None) == -6
    assert norm(16, 32, None, 0, 16, 1e-5, 1, 64, 0, None) == -3          # row stride shorter than the row
    emb = lib.pq_b200_embedding
    assert emb(None, 1, 16, 16, 16, 16, 64, 16, 64, 8, 8, 1, 0, None) == -6
    assert emb(16, 1, 16, 16, 16, 16, 32, 16, 64, 8, 8, 1, 0, None) == -3  # out stride shorter than k
    assert emb(16, 1, 16, 16, 16, 16, 64, 16, 64, 5, 8, 1, 0, None) in (-2, -3)   # d = 5
    assert emb(16, 0, 16, 16, 16, 16, 64, 16, 64, 8, 8, 1, 0, None) == 0   # no ids: nothing is launched
    two = lib.pq_b200_allreduce_twostage
    ptrs = (ctypes.c_void_p * 2)(16, 16)
    assert two(16, 16, 64, 7, ptrs, ptrs, ptrs, 0, 2, 1 << 20, None) == -1
    assert two(16, 16, 64, 0, ptrs, ptrs, ptrs, 2, 2, 1 << 20, None) == -3            # rank out of range
    assert two(16, 16, 60, 0, ptrs, ptrs, ptrs, 0, 2, 1 << 20, None) == -4            # not a multiple of 16 bytes
    assert two(16, 16, 1 << 20, 0, ptrs, ptrs, ptrs, 0, 2, 1 << 20, None) == -5       # larger than the buffers


def test_static_scratch_entry_point_rejects_without_gpu():
    """pq_b200_gemv_static (no reference counterpart): argument errors and the size limit are decided on the
    host, before any CUDA call."""
    lib = _lib.load()
    parts, n_parts = _lib.parts_array([4096])
    ok = (16, 4096, 16, 16, 16, None, 16, 4096)
    assert lib.pq_b200_gemv_static(*ok, 1, 4096, 4096, 8, 8, 1, parts, n_parts, 7, None) == -1        # dtype
    assert lib.pq_b200_gemv_static(*ok, 1, 4096, 4096, 8, 17, 1, parts, n_parts, 0, None) == -2       # format
    assert lib.pq_b200_gemv_static(*ok, 1, 4096, 4095, 8, 8, 1, parts, n_parts, 0, None) == -3        # k % d
    assert lib.pq_b200_gemv_static(None, 4096, 16, 16, 16, None, 16, 4096, 1, 4096, 4096, 8, 8, 1, parts, n_parts, 0, None) == -6
    assert lib.pq_b200_gemv_static(16, 4096, 16, 24, 16, None, 16, 4096, 1, 4096, 4096, 8, 8, 1, parts, n_parts, 0, None) == -4
    assert lib.pq_b200_gemv_static(*ok, 9, 4096, 4096, 8, 8, 1, parts, n_parts, 0, None) == -7        # 9 tokens
    assert lib.pq_b200_gemv_static(*ok, 1, 4096, 4096, 8, 16, 1, parts, n_parts, 0, None) == -7       # 16-bit codes
    big, n_big = _lib.parts_array([28672, 28672])
    assert lib.pq_b200_workspace_bytes(_lib.ALGO_GEMV_LUT_TMA, 8, 57344, 8192, 8, 8, 1, 2) > (4 << 20)
    rc = lib.pq_b200_gemv_static(16, 8192, 16, 16, 16, None, 16, 57344, 8, 57344, 8192, 8, 8, 1, big, n_big, 0, None)
    assert rc == -9 and b"PQ_STATIC_SCRATCH_BYTES" in lib.pq_b200_strerror(rc)
    import re
    header = open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include", "pq_b200.h")).read()
    assert re.search(r"#define PQ_STATIC_SCRATCH_BYTES \(4 << 20\)", header)
