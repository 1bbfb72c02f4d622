"""Run one pq_gemm configuration a few times (for ncu captures).  Weights rotate over `--nsets`
distinct tensors so the codes are not L2-resident."""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import vllm_pq_b200 as P  # noqa: E402

ALGOS = {"auto": 0, "direct": 1, "lut": 2, "tc": 3, "mma": 4, "lut2": 5}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--parts", default="11008,11008")
    ap.add_argument("--k", type=int, default=4096)
    ap.add_argument("--fmt", default="8x8x1")
    ap.add_argument("--t", type=int, default=1)
    ap.add_argument("--algo", default="auto")
    ap.add_argument("--dtype", default="fp16")
    ap.add_argument("--nsets", type=int, default=4)
    ap.add_argument("--iters", type=int, default=8)
    ap.add_argument("--pdl", type=int, default=0)
    a = ap.parse_args()
    parts = [int(v) for v in a.parts.split(",")]
    b, d, m = (int(v) for v in a.fmt.split("x"))
    dtype = torch.float16 if a.dtype == "fp16" else torch.bfloat16
    n, e = sum(parts), 1 << b
    P.ops.set_pdl(bool(a.pdl))
    sets = []
    for _ in range(a.nsets):
        codes = torch.randint(-(e // 2), e // 2, (n, a.k // d, m), device="cuda", dtype=torch.int32).to(
            torch.int8 if b <= 8 else torch.int16)
        cb = torch.randn(len(parts) * m, e, 1, d, device="cuda", dtype=dtype)
        sc = (torch.rand(n, 1, 1, 1, device="cuda") + 0.5).to(dtype)
        sets.append((codes, cb, sc))
    x = torch.randn(a.t, a.k, device="cuda", dtype=dtype)
    for i in range(a.iters):
        c, cb, sc = sets[i % a.nsets]
        y = P.ops.pq_gemm(x, c, cb, sc, parts, None, algo=ALGOS[a.algo])
    torch.cuda.synchronize()
    print("ok", tuple(y.shape), float(y.float().abs().mean()))


if __name__ == "__main__":
    main()
