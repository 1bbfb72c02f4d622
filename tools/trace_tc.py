"""Phase stamps of the tcgen05 GEMM (a -DPQ_TC_TRACE build of the library, PQ_B200_LIB=...): ns from the
first CTA's start for CTA 0 and the median CTA.  Diagnostics only."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import vllm_pq_b200 as P  # noqa: E402
from vllm_pq_b200 import _lib  # noqa: E402

NAMES = ["start", "tmem alloc + barriers", "first A stage full (MMA warp)", "first B stage full", "last MMA committed",
         "epilogue: accumulator ready", "epilogue done", "codebooks staged", "first A stage written", "all warps done"]


def main():
    n, k, t = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
    b, d, m = (int(v) for v in (sys.argv[4] if len(sys.argv) > 4 else "8x8x1").split("x"))
    lib = _lib.load()
    e = 1 << b
    sets = []
    for _ in range(6):
        codes = torch.randint(-(e // 2), e // 2, (n, k // d, m), device="cuda", dtype=torch.int32).to(
            torch.int8 if b <= 8 else torch.int16)
        cb = torch.randn(m, e, 1, d, device="cuda", dtype=torch.float16)
        sc = (torch.rand(n, 1, 1, 1, device="cuda") + 0.5).half()
        sets.append((codes, cb, sc))
    x = torch.randn(t, k, device="cuda", dtype=torch.float16)
    for s in sets:
        P.ops.pq_gemm(x, s[0], s[1], s[2], [n], None, algo=_lib.ALGO_GEMM_TC)
    torch.cuda.synchronize()
    trace = torch.zeros(148 * 16, dtype=torch.int64, device="cuda")
    lib.pq_b200_debug_set_trace(trace.data_ptr())
    P.ops.pq_gemm(x, sets[0][0], sets[0][1], sets[0][2], [n], None, algo=_lib.ALGO_GEMM_TC)
    torch.cuda.synchronize()
    lib.pq_b200_debug_set_trace(None)
    tr = trace.view(148, 16).cpu()
    used = tr[:, 0] > 0
    tr = tr[used]
    t0 = int(tr[:, 0].min())
    print(f"n={n} k={k} t={t} {b}x{d}x{m}: {int(used.sum())} CTAs")
    for i, name in enumerate(NAMES):
        col = (tr[:, i] - t0).float()
        col = col[tr[:, i] > 0]
        if len(col):
            print(f"  {name:34s} first {col.min():8.0f}  median {col.median():8.0f}  last {col.max():8.0f} ns")


if __name__ == "__main__":
    main()
