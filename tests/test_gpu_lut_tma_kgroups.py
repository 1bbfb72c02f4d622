"""Cross-cluster split-K of the TMA-fed table GEMV (pq_gemv_lut2.cu, KG > 1): the partial sums of
a row block's K groups meet in the workspace and the last CTA to take a ticket adds them.

The planner picks (cluster size, tiles per CTA, K groups) from a cost model; PQ_LUT2_CS / PQ_LUT2_KG
force them.  They are read once per process, so every forced combination runs in a subprocess that
checks the kernel against the CPU oracle (fp64 contraction) and for bit-reproducibility, including
the case where the last K group leaves some ranks of its cluster without a tile."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

SCRIPT = r"""
import sys, torch
sys.path.insert(0, %(root)r)
import vllm_pq_b200 as P
from vllm_pq_b200 import _lib
from oracle import pq_oracle as O

ALGO = _lib.ALGO_GEMV_LUT_TMA
bad = 0
for (parts, groups, b, d, m, t, dtype) in [
        ([4096], 1376, 8, 8, 1, 1, torch.float16),          # llama-2-7b down_proj: 22 tiles
        ([1024, 512, 512], 256, 8, 8, 2, 1, torch.float16),
        ([200, 72], 144, 8, 8, 1, 3, torch.bfloat16),
        ([640], 208, 6, 8, 2, 2, torch.float16),            # 6.5 tiles: ragged last tile
        ([300], 96, 8, 4, 4, 1, torch.float16)]:
    k, n = groups * d, sum(parts)
    if not _lib.load().pq_b200_algo_supported(ALGO, t, n, k, d, b, m):
        continue
    c = O.make_case(parts, k, b, d, m, t=t, dtype=dtype, seed=3, bias=True)
    g = {kk: (v.cuda() if torch.is_tensor(v) else v) for kk, v in c.items()}
    y = P.ops.pq_gemm(g["x"], g["codes"], g["codebooks"], g["scales"], parts, g["bias"], algo=ALGO)
    for _ in range(3):      # the tickets must be back at zero after every launch
        y2 = P.ops.pq_gemm(g["x"], g["codes"], g["codebooks"], g["scales"], parts, g["bias"], algo=ALGO)
        assert torch.equal(y, y2), "not reproducible"
    exact = O.linear_fp32(c["x"], c["codes"], c["codebooks"], c["scales"], parts, c["bias"])
    err = O.error_report(y.cpu(), exact)
    tol = 2e-3 if dtype == torch.float16 else 1.6e-2      # the P3 bound of tests/test_gpu_parity.py
    print(parts, groups, (b, d, m), t, err)
    if not err["rms_rel"] <= tol:
        bad += 1
sys.exit(1 if bad else 0)
"""


@pytest.mark.parametrize("cs,kg", [(0, 0), (1, 2), (2, 2), (4, 3), (8, 2), (8, 3), (2, 8), (1, 5)])
def test_lut_tma_forced_k_groups(cs, kg):
    env = dict(os.environ)
    if cs:
        env["PQ_LUT2_CS"] = str(cs)
    if kg:
        env["PQ_LUT2_KG"] = str(kg)
    r = subprocess.run([sys.executable, "-c", SCRIPT % {"root": ROOT}], env=env, capture_output=True,
                       text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
