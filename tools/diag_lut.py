import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import vllm_pq_b200 as P
from vllm_pq_b200 import _lib
from oracle import pq_oracle as O
for t in (1, 2, 4):
    for parts, k in (([4096]*3, 4096), ([4096], 4096), ([1024], 4096), ([768], 4096), ([128], 4096), ([4096], 512)):
        c = O.make_case(parts, k, 8, 8, 1, t=t, dtype=torch.float16, seed=1)
        g = {kk: (v.cuda() if torch.is_tensor(v) else v) for kk, v in c.items()}
        outs = {}
        for name, algo in (("direct", 1), ("lut", 2)):
            y = P.ops.pq_gemm(g["x"], g["codes"], g["codebooks"], g["scales"], parts, None, algo=algo)
            y_again = P.ops.pq_gemm(g["x"], g["codes"], g["codebooks"], g["scales"], parts, None, algo=algo)
            y2 = P.ops.pq_gemm(g["x"] * 2, g["codes"], g["codebooks"], g["scales"], parts, None, algo=algo)
            torch.cuda.synchronize()
            nd = int((y != y_again).sum()); nl = int((y2 != y * 2).sum())
            md = float((y2.float() - 2 * y.float()).abs().max())
            outs[name] = y
            print(f"t={t} parts={parts} k={k} {name}: nondeterministic={nd} nonlinear={nl} maxdiff={md:.4g} nan={int(torch.isnan(y).sum())}")
        d = (outs["direct"].float() - outs["lut"].float()).abs()
        print("   direct-vs-lut max", float(d.max()), "rel", float(d.max() / outs["direct"].float().abs().max()))
