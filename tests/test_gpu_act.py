"""SiLU(gate) * up taken on the x load of down_proj (pq_b200_gemm_silu_mul)
and as its own launch (pq_b200_silu_and_mul), against the reference's SiluAndMul.forward_native
(activation.py:73-76) followed by the oracle's PQ linear."""
import pytest
import torch

import vllm_pq_b200 as P
from oracle import pq_oracle as O
from vllm_pq_b200 import _lib

pytestmark = pytest.mark.gpu

DTYPES = [torch.float16, torch.bfloat16]
ULP = {torch.float16: 2.0 ** -10, torch.bfloat16: 2.0 ** -7}


@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("t,k", [(1, 11008), (3, 344), (8, 4096), (64, 14336), (5, 6), (2, 1376)])
def test_silu_and_mul_matches_native(dtype, t, k):
    g = torch.Generator().manual_seed(t * 131 + k)
    x = (torch.randn(t, 2 * k, generator=g) * 2.0).to(dtype)
    ref = O.silu_and_mul(x)
    out = P.ops.silu_and_mul(x.cuda()).cpu()
    assert out.shape == ref.shape and out.dtype == dtype
    # identical rounding steps; expf may differ by one fp32 ulp, which moves a result by at most
    # one ulp of the dtype, and rarely
    diff = (out.float() - ref.float()).abs()
    assert bool((diff <= ULP[dtype] * ref.float().abs() + 1e-30).all())
    assert float((out.view(torch.int16) == ref.view(torch.int16)).float().mean()) > 0.995
    # strided input (a slice of a wider buffer)
    wide = torch.zeros(t, 2 * k + 8, dtype=dtype)
    wide[:, :2 * k] = x
    out2 = P.ops.silu_and_mul(wide.cuda()[:, :2 * k]).cpu()
    assert torch.equal(out2.view(torch.int16), out.view(torch.int16))


CASES = [  # parts, k, (b, d, m), t
    ([4096], 11008, (8, 8, 1), 1),            # llama-2-7b down_proj
    ([4096], 2752, (8, 8, 1), 1),             # its TP=4 shard: code pitch 344, cp.async loader
    ([1024], 1376, (8, 8, 2), 3),
    ([512, 256], 1024, (6, 8, 1), 8),
    ([640], 2048, (8, 4, 4), 2),
    ([768], 1024, (8, 16, 1), 5),
    ([1024], 2048, (16, 8, 1), 1),            # 16-bit codes: no table kernel, activation as its own launch
    ([512], 1024, (8, 8, 1), 48),             # prefill-sized: tensor-core kernels
]


@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("parts,k,fmt,t", CASES)
def test_gemm_silu_mul(dtype, parts, k, fmt, t):
    b, d, m = fmt
    c = O.make_case(parts, k, b, d, m, t=t, dtype=dtype, seed=11, bias=True)
    g = torch.Generator().manual_seed(5)
    x2 = (torch.randn(t, 2 * k, generator=g) * 1.5).to(dtype)
    act = O.silu_and_mul(x2)
    exact = O.linear_fp32(act, c["codes"], c["codebooks"], c["scales"], parts, c["bias"])
    dev = {kk: (v.cuda() if torch.is_tensor(v) else v) for kk, v in c.items()}
    tol = 2e-3 if dtype == torch.float16 else 1.6e-2          # the P3 bound of tests/test_gpu_parity.py
    lib = _lib.load()
    n = sum(parts)
    algos = [_lib.ALGO_AUTO] + [a for a in (_lib.ALGO_GEMV_LUT_TMA, _lib.ALGO_GEMV_DIRECT, _lib.ALGO_GEMM_MMA,
                                            _lib.ALGO_GEMM_TC)
                                if lib.pq_b200_algo_supported(a, t, n, k, d, b, m)]
    outs = {}
    for algo in algos:
        y = P.ops.pq_gemm_silu_mul(x2.cuda(), dev["codes"], dev["codebooks"], dev["scales"], parts, dev["bias"],
                                   algo=algo)
        err = O.error_report(y.cpu(), exact)
        assert err["rms_rel"] <= tol, (algo, err)
        outs[algo] = y
    # the fused x load and the separate launch see the same activations: same result as the plain
    # call of the same kernel on them (bit for bit - the activation values are identical)
    act_dev = P.ops.silu_and_mul(x2.cuda())
    for algo in algos:
        y_plain = P.ops.pq_gemm(act_dev, dev["codes"], dev["codebooks"], dev["scales"], parts, dev["bias"], algo=algo)
        assert torch.equal(outs[algo], y_plain), algo


def test_gemm_silu_mul_rejects_plain_width():
    c = O.make_case([256], 512, 8, 8, 1, t=1, dtype=torch.float16, seed=1)
    dev = {kk: (v.cuda() if torch.is_tensor(v) else v) for kk, v in c.items()}
    with pytest.raises(RuntimeError, match="gate\\|up"):
        P.ops.pq_gemm_silu_mul(dev["x"], dev["codes"], dev["codebooks"], dev["scales"], [256], None)
