"""SiLU(gate)*up + down_proj at decode: fused x load against silu_and_mul launch + plain GEMV.
Chains of 32 (gate_up -> [act] -> down) pairs in one CUDA graph with programmatic dependent launch,
like bench.py; prints microseconds per pair.  Usage: python tools/bench_act.py [--shape 7b|70b]"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import vllm_pq_b200 as P  # noqa: E402
from vllm_pq_b200 import stack as S  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shape", default="llama-2-7b")
    ap.add_argument("--pairs", type=int, default=32)
    ap.add_argument("--iters", type=int, default=30)
    args = ap.parse_args()
    shape = S.SHAPES[args.shape]
    dev = torch.device("cuda:0")
    P.ops.set_pdl(True)
    cfg = P.PQConfig(8, 8, 1, 1)
    with torch.device(dev):
        model = S.PQLlamaLinearStack(shape, cfg, torch.float16, n_layers=args.pairs).init_synthetic(seed=0)
    x = torch.randn(1, shape.hidden, dtype=torch.float16, device=dev)

    def run(mode):
        h = x
        for layer in model.layers:
            gu, _ = layer.gate_up_proj(h)
            if mode == "fused":
                h, _ = layer.down_proj(gu, input_act="silu_and_mul")
            elif mode == "separate":
                h, _ = layer.down_proj(P.ops.silu_and_mul(gu))
            else:                                   # no activation math at all (bench.py's stand-in)
                h, _ = layer.down_proj(gu[..., :layer.mlp_width])
        return h

    out = {}
    stream = torch.cuda.Stream()
    for mode in ("slice", "fused", "separate"):
        with torch.cuda.stream(stream):
            for _ in range(3):
                run(mode)
            stream.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=stream):
                y = run(mode)
            for _ in range(5):
                g.replay()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            stream.synchronize()
            e0.record(stream)
            for _ in range(args.iters):
                g.replay()
            e1.record(stream)
            stream.synchronize()
        out[mode] = e0.elapsed_time(e1) * 1e3 / (args.iters * args.pairs)
        out[mode + "_finite"] = bool(torch.isfinite(y).all())
    out["shape"] = args.shape
    out["unit"] = "us per gate_up + down pair, one token, 8x8x1 fp16, graph + PDL"
    print(json.dumps(out))


if __name__ == "__main__":
    main()
