"""Absolute phase timeline of a CHAIN of table-GEMV launches (programmatic dependent launch): every
launch stamps %globaltimer into its own buffer; times are printed relative to the first CTA start of
the first traced launch, so the overlap (or not) of launch i+1's prologue with launch i is visible."""
import argparse
import ctypes
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import vllm_pq_b200 as P  # noqa: E402
from vllm_pq_b200 import _lib  # noqa: E402

PH = {0: "start", 8: "barriers ready", 1: "consts done", 2: "pdl wait done", 3: "x staged", 4: "table 0 built",
      5: "lookups done", 6: "cluster barrier", 7: "y stored"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=4096)
    ap.add_argument("--k", type=int, default=4096)
    ap.add_argument("--parts", type=int, default=1)
    ap.add_argument("--pdl", type=int, default=1)
    ap.add_argument("--chain", type=int, default=4)
    ap.add_argument("--nsets", type=int, default=40)
    a = ap.parse_args()
    lib = _lib.load()
    P.ops.set_pdl(bool(a.pdl))
    parts = [a.n // a.parts] * a.parts
    sets = []
    for _ in range(a.nsets):
        codes = torch.randint(-128, 128, (a.n, a.k // 8, 1), device="cuda", dtype=torch.int32).to(torch.int8)
        cb = torch.randn(a.parts, 256, 1, 8, device="cuda", dtype=torch.float16)
        sc = (torch.rand(a.n, 1, 1, 1, device="cuda") + 0.5).half()
        sets.append((codes, cb, sc))
    x = torch.randn(1, a.k, device="cuda", dtype=torch.float16)
    traces = [torch.zeros(1 << 14, dtype=torch.int64, device="cuda") for _ in range(a.chain)]
    algo = _lib.ALGO_GEMV_LUT_TMA
    for _ in range(3):
        for s in sets:
            P.ops.pq_gemm(x, s[0], s[1], s[2], parts, None, algo=algo)
    torch.cuda.synchronize()
    # one CUDA graph: 4 untraced lead-in launches, then the traced ones (the trace pointer is a kernel
    # argument fixed at capture time); eager launches from Python are ~40 us apart and show nothing
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        for s in sets[:2]:                                   # the per-stream workspace is born outside the capture
            P.ops.pq_gemm(x, s[0], s[1], s[2], parts, None, algo=algo)
        st.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=st):
            keep = []
            for s in sets[a.chain:a.chain + 4]:
                keep.append(P.ops.pq_gemm(x, s[0], s[1], s[2], parts, None, algo=algo))
            for i in range(a.chain):
                lib.pq_b200_debug_set_trace(traces[i].data_ptr())
                s = sets[i]
                keep.append(P.ops.pq_gemm(x, s[0], s[1], s[2], parts, None, algo=algo))
            lib.pq_b200_debug_set_trace(None)
        for _ in range(3):
            graph.replay()
    torch.cuda.synchronize()
    g = (ctypes.c_int * 4)()
    lib.pq_b200_debug_last_launch(g)
    n_cta = g[0] * g[1]
    print(f"grid=({g[0]},{g[1]}) cluster={g[2]} smem={g[3]} B  pdl={a.pdl}")
    t0 = None
    for i in range(a.chain):
        tr = traces[i][: n_cta * 16].view(n_cta, 16).cpu()
        if t0 is None:
            t0 = int(tr[:, 0].min())
        row = []
        for slot, name in PH.items():
            col = (tr[:, slot] - t0).float()
            row.append(f"{name} {col.min():.0f}/{col.median():.0f}/{col.max():.0f}")
        print(f"launch {i}: " + " | ".join(row))


if __name__ == "__main__":
    main()
