"""Load tests/golden/*.npz (written by tests/golden/make_golden.py from the reference's own code)."""
import glob
import os

import numpy as np
import torch

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def structured_codebook(n_books, nbits, d, dtype):
    """Must stay identical to tests/golden/make_golden.py:structured_codebook."""
    e = 1 << nbits
    book = torch.arange(n_books).view(-1, 1, 1, 1)
    ent = torch.arange(e).view(1, -1, 1, 1)
    j = torch.arange(d).view(1, 1, 1, -1)
    v = ((ent * 37 + j * 101 + book * 53 + (ent >> 5) * 7) % 256 - 128).float() / 8.0
    return v.to(dtype)


def names():
    return sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN_DIR, "*.npz")))


def load(name):
    z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    b, d, m = (int(v) for v in z["fmt"])
    dtype = torch.bfloat16 if bool(z["is_bf16"]) else torch.float16
    parts = [int(v) for v in z["partition_sizes"]]

    def f(key):
        return torch.from_numpy(z[key].copy()).view(dtype)

    if bool(z["structured"]):
        codebooks = structured_codebook(len(parts) * m, b, d, dtype)
    else:
        codebooks = f("codebooks")
    return dict(name=name, nbits=b, d=d, m=m, dtype=dtype, partition_sizes=parts,
                codes=torch.from_numpy(z["codes"].copy()), codebooks=codebooks, scales=f("scales"),
                x=f("x"), bias=f("bias") if "bias" in z.files else None,
                w_unscaled=f("w_unscaled"), w_scaled=f("w_scaled"), y=f("y"))


def same_bits(a, b):
    return torch.equal(a.contiguous().view(torch.int16).cpu(), b.contiguous().view(torch.int16).cpu())
