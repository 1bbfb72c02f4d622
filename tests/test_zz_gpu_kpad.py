"""K-padded private copy on the device (Llama-2-7B down_proj at TP = 8: K = 1376), through the layer API.

The kernels it reaches (mma.sync, tcgen05 at K' = 1408; the decode kernels over the checkpoint layout at
1..8 tokens) are covered at other shapes by the earlier files; the host logic is covered on the CPU in
tests/test_kpad_cpu.py.  Collected last: it was the last path added in round 2."""
import pytest
import torch

import vllm_pq_b200 as P
from oracle import pq_oracle as O
from tests.test_gpu_parity import check_output
from vllm_pq_b200 import _lib

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("fmt", [(8, 8, 1), (8, 8, 2), (16, 8, 1)], ids=["8x8x1", "8x8x2", "16x8x1"])
@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
def test_down_proj_tp8_shard_all_token_counts(fmt, dtype):
    b, d, m = fmt
    k, n = 1376, 4096
    c = O.make_case([n], k, b, d, m, t=300, dtype=dtype, seed=61)
    c["codebooks"] = (c["codebooks"].float() / (k * m) ** 0.5).to(dtype)
    with torch.device("cuda"):
        lin = P.RowParallelLinear(k, n, bias=False, params_dtype=dtype, quant_config=P.PQConfig(d, b, m, 1))
    lin.codes.data.copy_(c["codes"]); lin.codebooks.data.copy_(c["codebooks"]); lin.scales.data.copy_(c["scales"])
    x = c["x"].cuda()
    before = {t: lin(x[:t])[0] for t in (1, 9, 64)}              # checkpoint layout: gather GEMV above 8 tokens
    lin.quant_method.process_weights_after_loading(lin)
    gp = lin.codes_kpad.shape[1]
    assert gp * d == 1408 and torch.equal(lin.codes_kpad[:, :k // d], lin.codes.data) and not lin.codes_kpad[:, k // d:].any()
    if b == 8:
        assert lin.codes_kpad.data_ptr() == lin.codes_padded.data_ptr()      # one private copy, two views
    lib = _lib.load()
    assert lib.pq_b200_pick_algo(64, n, k, d, b, m) == _lib.ALGO_GEMV_DIRECT
    assert lib.pq_b200_pick_algo(64, n, gp * d, d, b, m) == _lib.ALGO_GEMM_TC
    for t in (1, 2, 8, 9, 16, 17, 64, 300):
        ct = dict(c); ct["x"] = c["x"][:t]
        y, _ = lin(x[:t])
        check_output(y, ct, f"t={t}")
        if t in before:     # same weights, another kernel family: equal within two linears' tolerance
            rel = float((y.double() - before[t].double()).pow(2).mean().sqrt() / before[t].double().pow(2).mean().sqrt())
            assert rel < (4e-3 if dtype == torch.float16 else 3.2e-2), (t, rel)
    # SiLU*mul on the input and the separate-launch form of the fused chain take the same copy
    gu = torch.randn(40, 2 * k, device="cuda", dtype=dtype)
    got, _ = lin(gu, input_act="silu_and_mul")
    ca = dict(c); ca["x"] = O.silu_and_mul(gu.cpu())
    check_output(got, ca, "silu_mul t=40")
    res = torch.randn(40, n, device="cuda", dtype=dtype)
    out, _ = lin(x[:40], residual=res)
    assert torch.equal(out, lin(x[:40])[0] + res)
    w = lin.quant_method.dequantize(lin, scaled=True)
    assert torch.equal(w.cpu().view(torch.int16),
                       O.dequantize_partitions(c["codes"], c["codebooks"], [n], c["scales"]).view(torch.int16))
