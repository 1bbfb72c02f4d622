"""Generate tests/golden/*.npz from the REFERENCE's own code (run in the build
container only; /root/reference does not exist on the GPU box).

    python tests/golden/make_golden.py

Each fixture holds the inputs and what the reference computed from them:
  w_unscaled  dequantize_weight(unpack_int_data(codes), codebooks_p, None)   aqlm.py:37-74
  w_scaled    the same with scales                                          aqlm.py:70-72
  y           generic_dequantize_gemm(x, codes, codebooks, scales, parts, bias)  aqlm.py:94-127
all per partition, concatenated, stored as raw 16-bit patterns.
The 16-bit-code case uses a formula codebook (exact in fp16) so that the 1 MiB
table need not be stored; ``structured_codebook`` is shared with the tests.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from oracle import pq_oracle as O          # noqa: E402
from oracle import ref_loader              # noqa: E402


def structured_codebook(n_books: int, nbits: int, d: int, dtype) -> torch.Tensor:
    """Entry values are multiples of 1/8 in [-16, 16): exact in fp16 and bf16."""
    e = 1 << nbits
    book = torch.arange(n_books).view(-1, 1, 1, 1)
    ent = torch.arange(e).view(1, -1, 1, 1)
    j = torch.arange(d).view(1, 1, 1, -1)
    v = ((ent * 37 + j * 101 + book * 53 + (ent >> 5) * 7) % 256 - 128).float() / 8.0
    return v.to(dtype)


# name, parts, K, (b, d, M), T, dtype, bias, structured codebook?
CASES = [
    ("c1x8_d8",      [48],         128, (8, 8, 1),  2, torch.float16, False, False),
    ("c2x8_qkv",     [32, 16, 16], 128, (8, 8, 2),  3, torch.float16, True,  False),
    ("c1x16_gateup", [24, 24],      64, (16, 8, 1), 2, torch.float16, False, True),
    ("c1x4_d4",      [32, 32],      64, (4, 4, 1),  1, torch.float16, False, False),
    ("c1x12_d16",    [40],         128, (12, 16, 1), 2, torch.float16, True, False),
    ("c3x8",         [40],          64, (8, 8, 3),  1, torch.float16, False, False),
    ("c8x8",         [24],          64, (8, 8, 8),  2, torch.float16, False, False),
    ("c1x8_ragged",  [37],      8 * 43, (8, 8, 1),  5, torch.float16, True,  False),
    ("c1x8_bf16",    [48],         128, (8, 8, 1),  2, torch.bfloat16, False, False),
    ("c1x12_d4",     [16, 8],       32, (12, 4, 1), 1, torch.float16, False, False),
]


def bits(t: torch.Tensor) -> np.ndarray:
    return t.contiguous().view(torch.int16).numpy()


def main():
    ref = ref_loader.load_reference_aqlm()
    for seed, (name, parts, k, (b, d, m), t, dtype, bias, structured) in enumerate(CASES):
        c = O.make_case(parts, k, b, d, m, t=t, dtype=dtype, seed=100 + seed, bias=bias)
        if structured:
            c["codebooks"] = structured_codebook(len(parts) * m, b, d, dtype)
        # force the signed-storage corner values into the first row
        e = 1 << b
        flat = c["codes"].view(-1)
        corner = torch.tensor([-1, -(e // 2), e // 2 - 1, 0], dtype=flat.dtype)
        flat[:4] = corner
        wu, ws, off = [], [], 0
        for p, n_p in enumerate(parts):
            cu = ref.unpack_int_data(c["codes"][off:off + n_p], b)
            cb = c["codebooks"][p * m:(p + 1) * m]
            wu.append(ref.dequantize_weight(cu, cb, None))
            ws.append(ref.dequantize_weight(cu, cb, c["scales"][off:off + n_p]))
            off += n_p
        y = ref.generic_dequantize_gemm(c["x"], c["codes"], c["codebooks"], c["scales"],
                                        parts, c["bias"])
        out = dict(
            codes=c["codes"].numpy(), scales=bits(c["scales"]), x=bits(c["x"]),
            partition_sizes=np.asarray(parts, np.int64),
            fmt=np.asarray([b, d, m], np.int64), is_bf16=np.asarray(dtype == torch.bfloat16),
            structured=np.asarray(structured),
            w_unscaled=bits(torch.cat(wu, 0)), w_scaled=bits(torch.cat(ws, 0)), y=bits(y))
        if not structured:
            out["codebooks"] = bits(c["codebooks"])
        if bias:
            out["bias"] = bits(c["bias"])
        path = os.path.join(HERE, name + ".npz")
        np.savez_compressed(path, **out)
        print(f"{name}: {os.path.getsize(path) / 1024:.1f} KiB")


if __name__ == "__main__":
    main()
