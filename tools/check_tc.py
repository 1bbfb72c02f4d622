"""Quick correctness probe of the tcgen05 GEMM against a dense fp32 GEMM over pq_dequant's weights."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import vllm_pq_b200 as P  # noqa: E402
from vllm_pq_b200 import _lib  # noqa: E402


def case(parts, k, b, d, m, t, dtype, bias=False, seed=0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    n, e = sum(parts), 1 << b
    codes = torch.randint(-(e // 2), e // 2, (n, k // d, m), device="cuda", generator=g, dtype=torch.int32).to(
        torch.int8 if b <= 8 else torch.int16)
    cb = torch.randn(len(parts) * m, e, 1, d, device="cuda", generator=g).to(dtype)
    sc = (torch.rand(n, 1, 1, 1, device="cuda", generator=g) + 0.5).to(dtype)
    x = torch.randn(t, k, device="cuda", generator=g).to(dtype)
    bi = torch.randn(n, device="cuda", generator=g).to(dtype) if bias else None
    return codes, cb, sc, x, bi


def run(parts, k, b, d, m, t, dtype, bias=False):
    codes, cb, sc, x, bi = case(parts, k, b, d, m, t, dtype, bias)
    y = P.ops.pq_gemm(x, codes, cb, sc, parts, bi, algo=_lib.ALGO_GEMM_TC)
    torch.cuda.synchronize()
    w = P.ops.pq_dequant(codes, cb, parts).double() * sc.double().view(-1, 1)
    ref = x.double() @ w.t()
    if bi is not None:
        ref = ref + bi.double()
    err = (y.double() - ref)
    rms = float(err.pow(2).mean().sqrt() / ref.pow(2).mean().sqrt())
    bad = int((~torch.isfinite(y)).sum())
    print(f"parts={parts} k={k} fmt={b}x{d}x{m} t={t} {str(dtype)[6:]} bias={bias}: rms_rel={rms:.2e} nonfinite={bad}",
          flush=True)
    return rms


if __name__ == "__main__":
    ok = True
    for args in [
        ([128], 64, 8, 8, 1, 16, torch.float16),
        ([128], 128, 8, 8, 1, 16, torch.float16),
        ([128], 512, 8, 8, 1, 16, torch.float16),
        ([256], 512, 8, 8, 1, 32, torch.bfloat16),
        ([128, 64, 72], 512, 8, 8, 2, 33, torch.float16, True),
        ([4096], 4096, 8, 8, 1, 256, torch.float16),
        ([4096], 4096, 16, 8, 1, 300, torch.bfloat16),
        ([1024, 1024], 1024, 4, 4, 1, 100, torch.float16),
        ([1024], 1024, 12, 16, 1, 64, torch.bfloat16),
        ([11008, 11008], 4096, 8, 8, 1, 2048, torch.float16),
    ]:
        r = run(*args)
        ok &= r < (2e-3 if args[6] == torch.float16 else 1.6e-2)
    print("ALL OK" if ok else "FAILED")
    sys.exit(0 if ok else 1)
