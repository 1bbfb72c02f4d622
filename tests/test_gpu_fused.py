"""The table GEMV's fused prologue / epilogue on one GPU: residual add, per-CTA sums of
squares, RMSNorm on the x load, and the stand-alone (residual add +) RMSNorm kernel - against the
oracle's restatement of csrc/layernorm_kernels.cu and against the unfused launches.  The all-reduce
part of the epilogue needs >= 2 GPUs: tools/tp_check.py (tests/test_gpu_tp.py)."""
import pytest
import torch

import vllm_pq_b200 as P
from oracle import pq_oracle as O
from tests.test_gpu_parity import RMS_REL_TOL, dev
from vllm_pq_b200 import _lib

pytestmark = pytest.mark.gpu
EPS = 1e-5


def ulp_close(a, b, dtype, frac_exact=0.98):
    """Same bits almost everywhere, never more than two ulps apart.  RMSNorm rounds twice (T(T(z * rstd) * w),
    layernorm_kernels.cu:91-97) and rstd comes from an fp32 reduction whose order differs from the oracle's
    fp64 sum: a last-bit difference in rstd moves the first rounding by one ulp on the few elements that sit
    on a rounding boundary, and the multiply by w in [0.5, 1.5) can turn that into two ulps of the result
    (seen once in some hundred runs on random data, gpurun r2_39)."""
    ai, bi = a.view(torch.int16).int(), b.view(torch.int16).int()
    diff = (ai - bi).abs()
    assert int(diff.max()) <= 2, int(diff.max())
    assert float((diff == 0).float().mean()) >= frac_exact, float((diff == 0).float().mean())


@pytest.fixture(autouse=True)
def _seeded():
    torch.manual_seed(1234)                                  # also seeds the CUDA generators: same data every run
    yield


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("shape", [(1, 4096), (5, 8192), (3, 1000)])
@pytest.mark.parametrize("with_residual", [True, False])
def test_fused_add_rms_norm_kernel(dtype, shape, with_residual):
    g = torch.Generator().manual_seed(3)
    x = torch.randn(shape, generator=g).to(dtype)
    r = torch.randn(shape, generator=g).to(dtype) if with_residual else None
    w = (torch.rand(shape[1], generator=g) + 0.5).to(dtype)
    out_ref, z_ref = O.fused_add_rms_norm(x, r, w, EPS)
    xd, wd = x.cuda(), w.cuda()
    rd = r.cuda() if with_residual else None
    P.ops.fused_add_rms_norm(xd, rd, wd, EPS)
    if with_residual:
        assert torch.equal(rd.cpu().view(torch.int16), z_ref.view(torch.int16)), "residual stream must be bit-exact"
    ulp_close(xd.cpu(), out_ref, dtype)


CASES = [("o7b", [4096], 4096, (8, 8, 1)), ("down7b", [4096], 11008, (8, 8, 1)), ("o7b_2x8", [4096], 4096, (8, 8, 2)),
         ("small", [512], 1024, (8, 8, 1)), ("d4", [1024], 2048, (8, 4, 1))]


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
@pytest.mark.parametrize("t", [1, 3, 8])
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
def test_residual_and_ssq_epilogue(case, t, dtype):
    _, parts, k, (b, d, m) = case
    c = O.make_case(parts, k, b, d, m, t=t, dtype=dtype, seed=21, bias=True)
    c["codebooks"] = (c["codebooks"].float() / (k * m) ** 0.5).to(dtype)
    g = dev(c)
    n = sum(parts)
    res = torch.randn(t, n, dtype=dtype, device="cuda")
    plain = P.ops.pq_gemm(g["x"], g["codes"], g["codebooks"], g["scales"], parts, g["bias"], algo=_lib.ALGO_GEMV_LUT_TMA)
    y, ssq = P.ops.pq_gemm_fused(g["x"], g["codes"], g["codebooks"], g["scales"], parts, g["bias"],
                                 residual=res, want_ssq=True)
    assert torch.equal(y, plain + res), "epilogue residual add != separate add (same roundings expected)"
    want = y.float().pow(2).sum(-1)
    got = ssq.sum(-1)
    assert ssq.dtype == torch.float32 and ssq.shape[0] == t
    assert torch.allclose(got, want, rtol=1e-5), (got, want)
    # no residual, no ssq: the fused entry point is the plain kernel
    y0, _ = P.ops.pq_gemm_fused(g["x"], g["codes"], g["codebooks"], g["scales"], parts, g["bias"])
    assert torch.equal(y0, plain)
    # reproducible
    y2, ssq2 = P.ops.pq_gemm_fused(g["x"], g["codes"], g["codebooks"], g["scales"], parts, g["bias"],
                                   residual=res, want_ssq=True)
    assert torch.equal(y, y2) and torch.equal(ssq, ssq2)


NORM_CASES = [("qkv7b", [4096] * 3, 4096, (8, 8, 1)), ("gateup7b", [11008] * 2, 4096, (8, 8, 1)),
              ("qkv_2x8", [1024, 256, 256], 2048, (8, 8, 2)), ("d16", [2048], 4096, (8, 16, 1))]


@pytest.mark.parametrize("case", NORM_CASES, ids=[c[0] for c in NORM_CASES])
@pytest.mark.parametrize("t", [1, 4])
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
def test_rms_norm_on_x_load(case, t, dtype):
    """producer (K -> hidden, residual add, ssq) -> consumer (RMSNorm on the x load) against the oracle:
    the consumer's input is the oracle's RMSNorm of the producer's z."""
    _, parts, hidden, (b, d, m) = case
    prod = O.make_case([hidden], 2048, 8, 8, 1, t=t, dtype=dtype, seed=31)
    prod["codebooks"] = (prod["codebooks"].float() / 2048 ** 0.5).to(dtype)
    gp = dev(prod)
    res = torch.randn(t, hidden, dtype=dtype, device="cuda")
    z, ssq = P.ops.pq_gemm_fused(gp["x"], gp["codes"], gp["codebooks"], gp["scales"], [hidden], None,
                                 residual=res, want_ssq=True)
    w = (torch.rand(hidden) + 0.5).to(dtype)
    c = O.make_case(parts, hidden, b, d, m, t=t, dtype=dtype, seed=32)
    g = dev(c)
    y, _ = P.ops.pq_gemm_fused(z, g["codes"], g["codebooks"], g["scales"], parts, None,
                               norm_ssq=ssq, norm_weight=w.cuda(), norm_eps=EPS)
    xn_ref, _ = O.fused_add_rms_norm(z.cpu(), None, w, EPS)
    ref = O.linear_fp32(xn_ref, c["codes"], c["codebooks"], c["scales"], parts, None)
    err = O.error_report(y.cpu(), ref)
    assert err["rms_rel"] < RMS_REL_TOL[dtype], err
    # against the separate launches: norm kernel, then the plain table GEMV (rstd may differ in its
    # last bit between the two reductions, so a few inputs differ by an ulp)
    xn = z.clone()
    P.ops.fused_add_rms_norm(xn, None, w.cuda(), EPS)
    ulp_close(xn.cpu(), xn_ref, dtype)
    y_sep = P.ops.pq_gemm(xn, g["codes"], g["codebooks"], g["scales"], parts, None, algo=_lib.ALGO_GEMV_LUT_TMA)
    err2 = O.error_report(y.cpu(), y_sep.double().cpu())
    assert err2["rms_rel"] < RMS_REL_TOL[dtype] / 2, err2


def test_layer_api_fused_chain_matches_separate_launches():
    """RowParallelLinear(residual, want_ssq) -> ColumnParallelLinear(input_norm) == the reference's chain
    apply -> (+ residual) -> RMSNorm -> apply, within tolerance; 16-token batch takes the unfused path."""
    dtype = torch.float16
    cfg = P.PQConfig(8, 8, 1, 1)
    kw = dict(bias=False, params_dtype=dtype, quant_config=cfg)
    with torch.device("cuda"):
        o = P.RowParallelLinear(4096, 4096, **kw)
        gu = P.MergedColumnParallelLinear(4096, [11008, 11008], **kw)
    gen = torch.Generator(device="cuda").manual_seed(5)
    for lin in (o, gu):
        lin.codes.data.copy_(torch.randint(-128, 128, lin.codes.shape, device="cuda", generator=gen, dtype=torch.int32))
        cb = torch.randn(lin.codebooks.shape, device="cuda", generator=gen) / 64
        lin.codebooks.data.copy_(cb - cb.mean(1, keepdim=True))
        lin.scales.data.copy_(torch.rand(lin.scales.shape, device="cuda", generator=gen) + 0.5)
    w = (torch.rand(4096, device="cuda") + 0.5).to(dtype)
    for t in (1, 4, 16):
        x = torch.randn(t, 4096, dtype=dtype, device="cuda")
        res = torch.randn(t, 4096, dtype=dtype, device="cuda")
        z, _, ssq = o(x, residual=res, want_ssq=True)
        y, _ = gu(z, input_norm=(ssq, w, EPS))
        # separate launches
        h, _ = o(x)
        z_ref = h + res
        xn = z_ref.clone()
        P.ops.fused_add_rms_norm(xn, None, w, EPS)
        y_ref, _ = gu(xn)
        if t == 4:        # separate launches use the mma.sync kernel at 2..16 tokens: same sums, other rounding
            assert float((z.float() - z_ref.float()).abs().max()) < 2e-2
        else:
            assert torch.equal(z, z_ref)
        rel = float((y.double() - y_ref.double()).pow(2).mean().sqrt() / y_ref.double().pow(2).mean().sqrt())
        assert rel < 1e-3, (t, rel)


@pytest.mark.parametrize("k", [1376, 2752, 8 * 100])
@pytest.mark.parametrize("fmt", [(8, 8, 1), (8, 8, 2), (4, 8, 1)])
def test_padded_codes_copy(k, fmt):
    """process_weights_after_loading pads code rows that are not a multiple of 16 bytes (TP = 4 / 8
    shards of down_proj) into a private copy for the TMA stream: same bits as the cp.async loader
    over the checkpoint layout, dequant untouched."""
    b, d, m = fmt
    dtype = torch.float16
    cfg = P.PQConfig(d, b, m, 1)
    c = O.make_case([4096], k, b, d, m, t=8, dtype=dtype, seed=41)
    c["codebooks"] = (c["codebooks"].float() / (k * m) ** 0.5).to(dtype)
    with torch.device("cuda"):
        lin = P.RowParallelLinear(k, 4096, bias=False, params_dtype=dtype, quant_config=cfg)
    lin.codes.data.copy_(c["codes"]); lin.codebooks.data.copy_(c["codebooks"]); lin.scales.data.copy_(c["scales"])
    row = (k // d) * m
    assert not hasattr(lin, "codes_padded")
    x = c["x"].cuda()
    before = {t: lin(x[:t])[0] for t in (1, 3, 8)}
    lin.quant_method.process_weights_after_loading(lin)
    if row % 16:
        assert lin.codes_padded.shape == (4096, (row + 15) // 16 * 16)
    else:
        assert not hasattr(lin, "codes_padded")
    for t in (1, 3, 8):
        y, _ = lin(x[:t])
        ct = dict(c); ct["x"] = c["x"][:t]
        ref = O.linear_fp32(ct["x"], c["codes"], c["codebooks"], c["scales"], [4096], None)
        assert O.error_report(y.cpu(), ref)["rms_rel"] < RMS_REL_TOL[dtype]
        if t == 1 and b == 8:         # same kernel, same order of sums: the loader does not change a bit
            assert torch.equal(y, before[t])
    w = lin.quant_method.dequantize(lin, scaled=True)
    w_ref = O.dequantize_partitions(c["codes"], c["codebooks"], [4096], c["scales"])
    assert torch.equal(w.cpu().view(torch.int16), w_ref.view(torch.int16))


# ---- PQ vocabulary layers -----------------------------------------------------------
@pytest.mark.parametrize("fmt", [(8, 8, 1), (8, 8, 2), (16, 8, 1), (4, 4, 1), (12, 16, 2)])
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
def test_embedding_rows_bit_exact(fmt, dtype):
    b, d, m = fmt
    vocab, hidden = 1000, 512
    cfg = P.PQConfig(d, b, m, 1, lm_head_quantized=True)
    c = O.make_case([vocab], hidden, b, d, m, dtype=dtype, seed=51)
    with torch.device("cuda"):
        emb = P.VocabParallelEmbedding(vocab, hidden, params_dtype=dtype, quant_config=cfg)
    for name in ("codes", "codebooks", "scales"):
        p = getattr(emb, name)
        p.weight_loader(p, c[name].cuda())
    ids = torch.tensor([[0, 999, 17, 17], [512, 1, 1000, 1023]], device="cuda")     # 1000, 1023: padding rows
    out = emb(ids)
    assert out.shape == (2, 4, hidden) and out.dtype == dtype
    w_ref = O.dequantize_partitions(c["codes"], c["codebooks"], [vocab], c["scales"])
    want = torch.zeros(2, 4, hidden, dtype=dtype)
    for i in range(2):
        for j in range(4):
            if int(ids[i, j]) < vocab:
                want[i, j] = w_ref[int(ids[i, j])]
    real = (ids < vocab).cpu()
    assert torch.equal(out.cpu()[real].view(torch.int16), want[real].view(torch.int16))
    assert float(out.cpu()[~real].abs().max()) == 0.0        # padding rows: zero scale (+-0 either way)
    # and it is the row gather of this library's own dequant
    w = emb.quant_method.dequantize(emb, scaled=True)
    assert torch.equal(out, w[ids])
    # out-of-range ids (other shards' tokens under TP) give zero rows
    z = P.ops.pq_embedding(torch.tensor([-1, 5000], device="cuda"), emb.codes, emb.codebooks, emb.scales)
    assert float(z.abs().max()) == 0.0


@pytest.mark.parametrize("t", [1, 5, 40])
def test_lm_head_logits(t):
    """Llama-2 vocabulary (32000 x 4096) as a PQ lm_head: logits against the oracle on a row sample."""
    from tests.test_gpu_configs import check_sampled
    cfg = P.PQConfig(8, 8, 1, 1, lm_head_quantized=True)
    c = O.make_case([32000], 4096, 8, 8, 1, t=t, dtype=torch.float16, seed=52)
    with torch.device("cuda"):
        head = P.ParallelLMHead(32000, 4096, params_dtype=torch.float16, quant_config=cfg)
    for name in ("codes", "codebooks", "scales"):
        p = getattr(head, name)
        p.weight_loader(p, c[name].cuda())
    logits = head.compute_logits(c["x"].cuda())
    assert logits.shape == (t, 32000)
    check_sampled(logits, c, "lm_head")


# ---- nibble-packed private copy of 4-bit codes (repack in process_weights_after_loading) --
@pytest.mark.parametrize("fmt", [(4, 8, 1), (4, 4, 1), (4, 8, 2), (4, 16, 2)], ids=["4x8x1", "4x4x1", "4x8x2", "4x16x2"])
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
def test_nibble_packed_copy(fmt, dtype):
    """Two 4-bit codes per byte + a 256-entry product codebook == the 4-bit layer: layer outputs against
    the oracle (decode: packed copy; 16 tokens: checkpoint tensors), dequant untouched and bit-exact, the
    M = 2 product codebook equal to the dequantised weights bit for bit."""
    b, d, m = fmt
    parts, k = [1024, 512], 2048
    cfg = P.PQConfig(d, b, m, 1)
    c = O.make_case(parts, k, b, d, m, t=16, dtype=dtype, seed=61)
    c["codebooks"] = (c["codebooks"].float() / (k * m) ** 0.5).to(dtype)
    with torch.device("cuda"):
        lin = P.MergedColumnParallelLinear(k, parts, bias=False, params_dtype=dtype, quant_config=cfg)
    lin.codes.data.copy_(c["codes"]); lin.codebooks.data.copy_(c["codebooks"]); lin.scales.data.copy_(c["scales"])
    x = c["x"].cuda()
    before = lin(x[:1])[0]
    lin.quant_method.process_weights_after_loading(lin)
    assert lin.codes_nibble.dtype == torch.int8 and lin.codebooks_pair.shape[1] == 256
    g = k // d
    assert tuple(lin.codes_nibble.shape) == ((sum(parts), g // 2, 1) if m == 1 else (sum(parts), g, 1))
    assert lin.codebooks_pair.shape[3] == (2 * d if m == 1 else d)
    for t in (1, 3, 8, 16):
        y, _ = lin(x[:t])
        ct = dict(c); ct["x"] = c["x"][:t]
        ref = O.linear_fp32(ct["x"], c["codes"], c["codebooks"], c["scales"], parts, None)
        err = O.error_report(y.cpu(), ref)
        assert err["rms_rel"] < RMS_REL_TOL[dtype], (t, err)
    rel = float((lin(x[:1])[0].double() - before.double()).pow(2).mean().sqrt() / before.double().pow(2).mean().sqrt())
    assert rel < RMS_REL_TOL[dtype], rel
    w = lin.quant_method.dequantize(lin, scaled=False)
    w_ref = O.dequantize_partitions(c["codes"], c["codebooks"], parts, None)
    assert torch.equal(w.cpu().view(torch.int16), w_ref.view(torch.int16))
    # the packed copy dequantises to the same weights through the same kernel (an 8-bit PQ layer)
    w_nib = P.ops.pq_dequant(lin.codes_nibble, lin.codebooks_pair, parts)
    assert torch.equal(w_nib, w)


# ---- per-expert PQ linears: the quantised Mixtral MoE block --------------------------------
@pytest.mark.parametrize("t", [1, 5, 40])
@pytest.mark.parametrize("fmt", [(8, 8, 1), (16, 8, 1)], ids=["8x8x1", "16x8x1"])
def test_mixtral_moe_block(t, fmt):
    """mixtral_quant.py:98-160 over PQ experts: against the same formula evaluated with the oracle."""
    import torch.nn.functional as F
    b, d, m = fmt
    dtype = torch.float16
    hidden, ffn, n_exp, top_k = 1024, 2048, 4, 2
    cfg = P.PQConfig(d, b, m, 1)
    with torch.device("cuda"):
        moe = P.PQMixtralMoE(n_exp, top_k, hidden, ffn, params_dtype=dtype, quant_config=cfg)
    gen = torch.Generator().manual_seed(71)
    cases = {}
    for e in range(n_exp):
        for name, n, k in (("w1", ffn, hidden), ("w3", ffn, hidden), ("w2", hidden, ffn)):
            c = O.make_case([n], k, b, d, m, dtype=dtype, seed=100 + 10 * e + len(name) + ord(name[1]))
            c["codebooks"] = (c["codebooks"].float() / (k * m) ** 0.5).to(dtype)
            cases[(e, name)] = c
            lin = getattr(moe.experts[e], name)
            lin.codes.data.copy_(c["codes"]); lin.codebooks.data.copy_(c["codebooks"]); lin.scales.data.copy_(c["scales"])
    gate = torch.randn(n_exp, hidden, generator=gen).to(dtype)
    moe.gate_weight.data.copy_(gate)
    x = torch.randn(t, hidden, generator=gen).to(dtype)
    y = moe(x.cuda())
    assert y.shape == (t, hidden)
    # the same formula on the CPU: routing in the layer's dtype as there, experts in fp64
    rw = F.softmax(F.linear(x.float(), gate.float()).to(dtype), dim=1, dtype=torch.float)
    rw, sel = torch.topk(rw, top_k, dim=-1)
    rw /= rw.sum(-1, keepdim=True)
    ref = torch.zeros(t, hidden, dtype=torch.float64)
    for e in range(n_exp):
        wgt = (rw * (sel == e)).sum(-1, keepdim=True).double()
        lin64 = lambda name, xin: O.linear_fp32(xin, cases[(e, name)]["codes"], cases[(e, name)]["codebooks"],   # noqa: E731
                                                cases[(e, name)]["scales"], [cases[(e, name)]["codes"].shape[0]], None)
        h = (F.silu(lin64("w1", x).to(dtype)) * lin64("w3", x).to(dtype))
        ref += lin64("w2", h) * wgt
    err = O.error_report(y.cpu(), ref)
    assert err["rms_rel"] < 4e-3, err      # three roundings (w1/w3 outputs, the product, the weighted sum) above one linear's bound
