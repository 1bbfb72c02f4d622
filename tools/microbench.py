"""Per-kernel timing on one GPU: rotates through enough distinct weight sets that codes do not
stay L2-resident (>= 256 MB per timed loop), CUDA events on the launching stream."""
import argparse
import ctypes
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import vllm_pq_b200 as P  # noqa: E402
from vllm_pq_b200 import _lib  # noqa: E402

HBM_GBS = 6551.4      # MEASURED_PEAKS.json (r2 pool)

try:
    import pynvml
    pynvml.nvmlInit()
    _H = pynvml.nvmlDeviceGetHandleByIndex(0)

    def sm_clock():
        return pynvml.nvmlDeviceGetClockInfo(_H, pynvml.NVML_CLOCK_SM)
except Exception:  # pragma: no cover
    def sm_clock():
        return -1

SHAPES = {
    "tiny": ([128], 512),
    "sq4096": ([4096], 4096),
    "qkv7b": ([4096] * 3, 4096),
    "o7b": ([4096], 4096),
    "gateup7b": ([11008] * 2, 4096),
    "down7b": ([4096], 11008),
    "gateup8b": ([14336] * 2, 4096),
    "down8b": ([4096], 14336),
    "gateup70b": ([28672] * 2, 8192),
    "down70b": ([8192], 28672),
    "down7b_tp4": ([4096], 2752),      # row-parallel shards: code pitch 344 / 172 bytes (not a multiple of 16)
    "down7b_tp8": ([4096], 1376),
    "down7b_tp2": ([4096], 5504),
    "n4096k1024": ([4096], 1024), "n4096k256": ([4096], 256), "n512k4096": ([512], 4096), "n128k256": ([128], 256),
}


def algo_bytes(t, n, k, d, b, m, p):
    sc = 1 if b <= 8 else 2
    return n * (k // d) * m * sc + p * m * (1 << b) * d * 2 + n * 2 + t * k * 2 + t * n * 2


def make(parts, k, b, d, m, dtype, nrep):
    n = sum(parts)
    e = 1 << b
    sets = []
    for _ in range(nrep):
        codes = torch.randint(-(e // 2), e // 2, (n, k // d, m), device="cuda", dtype=torch.int32).to(
            torch.int8 if b <= 8 else torch.int16)
        cb = torch.randn(len(parts) * m, e, 1, d, device="cuda", dtype=dtype)
        sc = (torch.rand(n, 1, 1, 1, device="cuda") + 0.5).to(dtype)
        sets.append((codes, cb, sc))
    return sets


def time_loop(fn, sets, iters, graph=True):
    """One CUDA graph holding one call per weight set; replayed `iters` times between events."""
    if graph:
        st = torch.cuda.Stream()
        with torch.cuda.stream(st):
            # warm up ON the capture stream: the op's workspace is per (device, stream) and is not
            # allocated during capture - without it the split-K plans are off (as in vLLM, which
            # also runs a model eagerly on the capture stream first)
            for s in sets[:2]:
                fn(s)
            st.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=st):
                keep = [fn(s) for s in sets]
        run = g.replay
    else:
        for s in sets[:2]:
            fn(s)
        torch.cuda.synchronize()

        def run():
            for s in sets:
                fn(s)
    import time
    t_end = time.time() + WARM_S          # let the SM clock ramp up before timing
    while time.time() < t_end:
        run()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        run()
    e1.record()
    clk = sm_clock()
    torch.cuda.synchronize()
    time_loop.last_clock = clk
    return e0.elapsed_time(e1) * 1e-3 / (iters * len(sets))


WARM_S = 0.3


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shapes", default="sq4096,qkv7b,gateup7b,down7b")
    ap.add_argument("--formats", default="8x8x1,8x8x2,16x8x1")
    ap.add_argument("--tokens", default="1,4")
    ap.add_argument("--algos", default="auto")
    ap.add_argument("--dtype", default="fp16")
    ap.add_argument("--pdl", type=int, default=0)
    ap.add_argument("--dequant", type=int, default=1)
    ap.add_argument("--out", default="")
    ap.add_argument("--graph", type=int, default=1)
    ap.add_argument("--warm", type=float, default=0.3)
    ap.add_argument("--nrep", type=int, default=0)
    ap.add_argument("--ref", type=int, default=0, help="also time the reference's own CUDA kernels (oracle/_ref)")
    ap.add_argument("--dense", type=int, default=0, help="also time torch.matmul on the dequantised dense weight (cuBLAS)")
    a = ap.parse_args()
    dtype = torch.float16 if a.dtype == "fp16" else torch.bfloat16
    P.ops.set_pdl(bool(a.pdl))
    global WARM_S
    WARM_S = a.warm
    names = {"auto": 0, "direct": 1, "lut": 2, "tc": 3, "mma": 4, "lut2": 5}
    rows = []
    for sname in a.shapes.split(","):
        parts, k = SHAPES[sname]
        n = sum(parts)
        for f in a.formats.split(","):
            b, d, m = (int(v) for v in f.split("x"))
            code_bytes = n * (k // d) * m * (1 if b <= 8 else 2)
            nrep = a.nrep or max(2, min(64, (256 << 20) // code_bytes + 1))
            sets = make(parts, k, b, d, m, dtype, nrep)
            if a.dequant:
                out_bytes = n * k * 2
                dt = time_loop(lambda s: P.ops.pq_dequant(s[0], s[1], parts), sets[:max(2, (512 << 20) // out_bytes)], 3)
                rows.append(dict(shape=sname, fmt=f, op="dequant", us=dt * 1e6,
                                 gbs=(out_bytes + code_bytes) / dt / 1e9))
                print(rows[-1], flush=True)
            for t in (int(v) for v in a.tokens.split(",")):
                x = torch.randn(t, k, device="cuda", dtype=dtype)
                for an in a.algos.split(","):
                    algo = names[an]
                    try:
                        dt = time_loop(lambda s: P.ops.pq_gemm(x, s[0], s[1], s[2], parts, None, algo=algo), sets, 20, graph=bool(a.graph))
                    except RuntimeError as exc:
                        print(f"skip {sname} {f} t={t} {an}: {exc}")
                        continue
                    ab = algo_bytes(t, n, k, d, b, m, len(parts))
                    plan = (ctypes.c_int * 8)()
                    _lib.load().pq_b200_debug_last_mma_plan(plan)
                    rows.append(dict(shape=sname, fmt=f, op="gemm", algo=an, t=t, us=dt * 1e6, mma_plan=list(plan),
                                     gbs=ab / dt / 1e9, frac=ab / dt / 1e9 / HBM_GBS,
                                     tflops=2 * t * n * k / dt / 1e12, sm_mhz=time_loop.last_clock))
                    print(rows[-1], flush=True)
                if a.dense:
                    dense = [(P.ops.pq_dequant(s[0], s[1], parts, s[2]),) for s in sets[:max(2, min(len(sets), (512 << 20) // (n * k * 2)))]]
                    dt = time_loop(lambda s: torch.matmul(x, s[0].t()), dense, 10, graph=bool(a.graph))
                    rows.append(dict(shape=sname, fmt=f, op="gemm", algo="dense_cublas", t=t, us=dt * 1e6,
                                     gbs=(n * k * 2) / dt / 1e9, tflops=2 * t * n * k / dt / 1e12, sm_mhz=time_loop.last_clock))
                    print(rows[-1], flush=True)
                    del dense
                if a.ref:
                    from oracle import ref_cuda as R
                    if R.available() and R.supports(b, d, m, dtype):
                        # what AQLMLinearMethod.apply runs: GEMV kernel up to 6 tokens, else dequant + F.linear
                        fn = R.aqlm_gemm if t <= 6 else R.optimized_dequantize_gemm
                        dt = time_loop(lambda s: fn(x, s[0], s[1], s[2], parts, None), sets, 5, graph=bool(a.graph))
                        ab = algo_bytes(t, n, k, d, b, m, len(parts))
                        rows.append(dict(shape=sname, fmt=f, op="gemm", algo="reference_cuda", t=t, us=dt * 1e6,
                                         gbs=ab / dt / 1e9, frac=ab / dt / 1e9 / HBM_GBS,
                                         tflops=2 * t * n * k / dt / 1e12, sm_mhz=time_loop.last_clock))
                        print(rows[-1], flush=True)
            del sets
            torch.cuda.empty_cache()
    if a.out:
        with open(a.out, "w") as fh:
            json.dump(rows, fh, indent=1)


if __name__ == "__main__":
    main()
