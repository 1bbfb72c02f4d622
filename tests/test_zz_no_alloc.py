"""pq_b200_gemv_static: the table GEMV with its K-group scratch in a static device array inside the library
(include/pq_b200.h; no counterpart in the reference).

Called straight through the C ABI (ctypes, raw pointers), compared with the existing entry point
pq_b200_gemm(PQ_ALGO_GEMV_LUT_TMA) given a caller workspace of the same size, within the tolerance the
parity tests use (rms-relative 2e-3 fp16 / 1.6e-2 bf16), and with the oracle on a sample of rows.  Two calls
in a row on one shape with a call on another shape in between: the array keeps the last call's contents.
Collected last (pytest -x must not lose other tests to a failure here); needs nothing
from the tests before it and leaves nothing behind - it passes alone and after them."""
import pytest
import torch

from oracle import pq_oracle as O
from tests.test_gpu_configs import check_sampled
from tests.test_gpu_parity import RMS_REL_TOL, dev
from vllm_pq_b200 import _lib

pytestmark = pytest.mark.gpu

STATIC_BYTES = 4 << 20                                  # PQ_STATIC_SCRATCH_BYTES
ERR_STATIC_SCRATCH = -9


def _common(g, x, y, parts, fmt):
    b, d, m = fmt
    arr, n_parts = _lib.parts_array(parts)
    t, k = x.shape
    n = sum(parts)
    dt = _lib.PQ_F16 if x.dtype == torch.float16 else _lib.PQ_BF16
    bias = None if g["bias"] is None else g["bias"].data_ptr()
    head = (x.data_ptr(), x.stride(0), g["codes"].data_ptr(), g["codebooks"].data_ptr(), g["scales"].data_ptr(),
            bias, y.data_ptr(), n, t, n, k, d, b, m, arr, n_parts, dt)
    return head, torch.cuda.current_stream().cuda_stream


def gemv_static(g, x, parts, fmt):
    y = torch.empty(x.shape[0], sum(parts), dtype=x.dtype, device="cuda")
    head, stream = _common(g, x, y, parts, fmt)
    rc = _lib.load().pq_b200_gemv_static(*head, stream)
    assert rc == 0, _lib.load().pq_b200_strerror(rc)
    return y


def gemv_workspace(g, x, parts, fmt, ws):
    y = torch.empty(x.shape[0], sum(parts), dtype=x.dtype, device="cuda")
    head, stream = _common(g, x, y, parts, fmt)
    rc = _lib.load().pq_b200_gemm(*head, _lib.ALGO_GEMV_LUT_TMA, ws.data_ptr(), ws.numel(), stream)
    assert rc == 0, _lib.load().pq_b200_strerror(rc)
    return y


def close(a, b, dtype):
    rel = float((a.double() - b.double()).pow(2).mean().sqrt() / b.double().pow(2).mean().sqrt())
    return rel <= RMS_REL_TOL[dtype]


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
def test_static_scratch_matches_workspace_entry_point(dtype):
    fmt = (8, 8, 1)
    ws = torch.zeros(STATIC_BYTES, dtype=torch.uint8, device="cuda")
    # down_proj of Llama-2-7B at one token (the planner splits K over CTA groups: partial sums and tickets
    # go through the scratch) and a fused qkv with bias at four tokens
    ca = O.make_case([4096], 11008, *fmt, t=1, dtype=dtype, seed=81)
    cb = O.make_case([1024, 1024, 1024], 4096, *fmt, t=4, dtype=dtype, seed=82, bias=True)
    ga, gb = dev(ca), dev(cb)
    assert _lib.load().pq_b200_workspace_bytes(_lib.ALGO_GEMV_LUT_TMA, 1, 4096, 11008, 8, 8, 1, 1) <= STATIC_BYTES
    want_a = gemv_workspace(ga, ga["x"], [4096], fmt, ws)
    want_b = gemv_workspace(gb, gb["x"], [1024, 1024, 1024], fmt, ws)
    got_a1 = gemv_static(ga, ga["x"], [4096], fmt)
    got_b = gemv_static(gb, gb["x"], [1024, 1024, 1024], fmt)          # other contents in the array ...
    got_a2 = gemv_static(ga, ga["x"], [4096], fmt)                    # ... before the second call on shape a
    got_a3 = gemv_static(ga, ga["x"], [4096], fmt)                    # and two calls back to back
    torch.cuda.synchronize()
    for got in (got_a1, got_a2, got_a3):
        assert torch.isfinite(got).all() and close(got, want_a, dtype)
    assert close(got_b, want_b, dtype)
    check_sampled(got_a2, ca, "static a")
    check_sampled(got_b, cb, "static b")


def test_static_scratch_rejects_what_does_not_fit():
    lib = _lib.load()
    fmt = (8, 8, 1)
    c = O.make_case([64], 64, *fmt, t=1, dtype=torch.float16, seed=83)       # pointers only: nothing is launched
    g = dev(c)
    n, k, t = 57344, 8192, 8                                                  # Llama-2-70B gate_up at 8 tokens: 14.7 MB
    assert lib.pq_b200_workspace_bytes(_lib.ALGO_GEMV_LUT_TMA, t, n, k, 8, 8, 1, 2) > STATIC_BYTES
    arr, n_parts = _lib.parts_array([n // 2, n // 2])
    rc = lib.pq_b200_gemv_static(g["x"].data_ptr(), k, g["codes"].data_ptr(), g["codebooks"].data_ptr(),
                                 g["scales"].data_ptr(), None, g["x"].data_ptr(), n, t, n, k, 8, 8, 1, arr, n_parts,
                                 _lib.PQ_F16, torch.cuda.current_stream().cuda_stream)
    assert rc == ERR_STATIC_SCRATCH and b"static" in lib.pq_b200_strerror(rc)
    torch.cuda.synchronize()
