"""Generate tests/golden/layers/*.npz from the REFERENCE's own layer classes (build container only).

    python tests/golden/make_golden_layers.py

The fused epilogue replaces the reference's RMSNorm / SiluAndMul launches; their
CPU form in the reference is ``forward_native`` (layers/layernorm.py:38-74, layers/activation.py:73-76).
Each fixture holds seeded inputs and what those methods returned, as raw 16-bit patterns:
  norm_*   x, residual, weight, eps ->  out          RMSNorm.forward_native(x)
                                        z, out_res   RMSNorm.forward_native(x, residual)   (z = new residual)
                                        out_z        RMSNorm.forward_native(z)
  silu_*   x -> out                                  SiluAndMul.forward_native(x)
``out_z`` is what the reference's CUDA kernel computes for (x, residual): it normalises the ROUNDED sum
(csrc/layernorm_kernels.cu:75-97) where forward_native carries the fp32 sum on (layernorm.py:45-48).
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from oracle import ref_loader              # noqa: E402

T, H = 6, 768
EPS = {"f16": 1e-5, "bf16": 1e-6}


def bits(t: torch.Tensor) -> np.ndarray:
    return t.contiguous().view(torch.int16).numpy()


def main():
    RMSNorm, SiluAndMul = ref_loader.load_reference_layers()
    out_dir = os.path.join(HERE, "layers")
    os.makedirs(out_dir, exist_ok=True)
    for seed, (tag, dtype) in enumerate((("f16", torch.float16), ("bf16", torch.bfloat16))):
        g = torch.Generator().manual_seed(500 + seed)
        x = (torch.randn(T, H, generator=g) * 1.5).to(dtype)
        r = (torch.randn(T, H, generator=g) * 3.0).to(dtype)
        w = (1.0 + 0.2 * torch.randn(H, generator=g)).to(dtype)
        norm = RMSNorm(H, EPS[tag])
        norm.weight.data = w
        with torch.no_grad():
            out = norm.forward_native(x)
            out_res, z = norm.forward_native(x, r)
            out_z = norm.forward_native(z)
        np.savez_compressed(os.path.join(out_dir, f"norm_{tag}.npz"), x=bits(x), residual=bits(r), weight=bits(w),
                            eps=np.asarray(EPS[tag]), is_bf16=np.asarray(dtype == torch.bfloat16),
                            out=bits(out), z=bits(z), out_res=bits(out_res), out_z=bits(out_z))
        gu = (torch.randn(T, 2 * H, generator=g) * 4.0).to(dtype)
        gu[0, :4] = torch.tensor([0.0, -0.0, 30.0, -30.0], dtype=dtype)
        act = SiluAndMul().forward_native(gu)
        np.savez_compressed(os.path.join(out_dir, f"silu_{tag}.npz"), x=bits(gu), out=bits(act),
                            is_bf16=np.asarray(dtype == torch.bfloat16))
    for f in sorted(os.listdir(out_dir)):
        print(f, f"{os.path.getsize(os.path.join(out_dir, f)) / 1024:.1f} KiB")


if __name__ == "__main__":
    main()
