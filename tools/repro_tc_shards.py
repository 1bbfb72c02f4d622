import sys, torch
sys.path.insert(0, '.')
import vllm_pq_b200 as P
from oracle import pq_oracle as O
from tests.test_gpu_configs import shards_70b, check_sampled, dev
for tp in (8, 4, 2):
    for name, parts, k in shards_70b(tp):
        c = O.make_case(parts, k, 8, 8, 1, t=2048, dtype=torch.float16, seed=3)
        g = dev(c)
        try:
            y = P.ops.pq_gemm(g["x"], g["codes"], g["codebooks"], g["scales"], parts, None)
            torch.cuda.synchronize()
            check_sampled(y, c, name)
            print("ok", tp, name, parts, k)
        except Exception as e:
            print("FAIL", tp, name, parts, k, str(e)[:200]); sys.exit(1)
