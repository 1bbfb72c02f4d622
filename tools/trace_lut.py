"""Phase timeline of the table GEMV (pq_gemv_lut.cu): every CTA stamps %globaltimer at its phase
boundaries into a debug buffer; this prints, per phase, when the first / median / last CTA got there
(ns from the first CTA's start).  Also prints cudaOccupancyMaxActiveClusters per cluster size."""
import argparse
import ctypes
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import vllm_pq_b200 as P  # noqa: E402
from vllm_pq_b200 import _lib  # noqa: E402

PHASES = ["start", "prefetch issued", "pdl wait done", "x staged + codebooks", "table 0 built",
          "lookups done", "pushed + cluster barrier", "reduced + y stored", "(lut_tma) barriers ready",
          "(lut_tma) item 0 done", "(lut_tma) warp 0 table stores issued", "(lut_tma) x loads issued",
          "(lut_tma) codebooks landed", "(lut_tma) table 1 built", "(lut_tma) tile 1 item 0 done",
          "(lut_tma) tile 0 items done"]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=22016)
    ap.add_argument("--k", type=int, default=4096)
    ap.add_argument("--parts", type=int, default=2)
    ap.add_argument("--m", type=int, default=1)
    ap.add_argument("--t", type=int, default=1)
    ap.add_argument("--nsets", type=int, default=12)
    ap.add_argument("--algo", type=int, default=_lib.ALGO_GEMV_LUT_TMA)
    ap.add_argument("--pdl", type=int, default=0)
    ap.add_argument("--chain", type=int, default=0, help="launch the traced kernel right after N others (PDL chain)")
    a = ap.parse_args()
    lib = _lib.load()
    P.ops.set_pdl(bool(a.pdl))
    for cw, smem in ((8, 100 * 1024), (8, 112 * 1024), (16, 200 * 1024)):
        print("max active clusters, %d consumer warps @%d KiB:" % (cw, smem // 1024),
              {cs: lib.pq_b200_debug_lut_tma_slots(cw, cs, smem) for cs in (1, 2, 4, 8, 16)})
    parts = [a.n // a.parts] * a.parts
    sets = []
    for _ in range(a.nsets):
        codes = torch.randint(-128, 128, (a.n, a.k // 8, a.m), device="cuda", dtype=torch.int32).to(torch.int8)
        cb = torch.randn(a.parts * a.m, 256, 1, 8, device="cuda", dtype=torch.float16)
        sc = (torch.rand(a.n, 1, 1, 1, device="cuda") + 0.5).half()
        sets.append((codes, cb, sc))
    x = torch.randn(a.t, a.k, device="cuda", dtype=torch.float16)
    trace = torch.zeros(1 << 16, dtype=torch.int64, device="cuda")
    for s in sets:          # warm-up (also evicts the first sets from L2)
        P.ops.pq_gemm(x, s[0], s[1], s[2], parts, None, algo=a.algo)
    torch.cuda.synchronize()
    for s in sets[1:1 + a.chain]:
        P.ops.pq_gemm(x, s[0], s[1], s[2], parts, None, algo=a.algo)
    lib.pq_b200_debug_set_trace(trace.data_ptr())
    P.ops.pq_gemm(x, sets[0][0], sets[0][1], sets[0][2], parts, None, algo=a.algo)
    torch.cuda.synchronize()
    lib.pq_b200_debug_set_trace(None)
    g = (ctypes.c_int * 4)()
    lib.pq_b200_debug_last_launch(g)
    n_cta = g[0] * g[1]
    print(f"grid=({g[0]},{g[1]}) cluster={g[2]} smem={g[3]} B")
    tr = trace[: n_cta * 16].view(n_cta, 16).cpu()
    t0 = int(tr[:, 0].min())
    for i, name in enumerate(PHASES):
        col = (tr[:, i] - t0).float()
        print(f"{name:24s} first {col.min():8.0f}  median {col.median():8.0f}  last {col.max():8.0f} ns")


if __name__ == "__main__":
    main()
