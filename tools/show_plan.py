"""Print the table GEMV's launch plan (pq_b200_debug_lut_tma_plan) for the Llama projections; runs
without a GPU (B200 occupancy figures are built in)."""
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from vllm_pq_b200 import _lib  # noqa: E402

SHAPES = {
    "7b": [("qkv", [4096] * 3, 4096), ("o", [4096], 4096), ("gate_up", [11008] * 2, 4096), ("down", [4096], 11008)],
    "70b": [("qkv", [8192, 1024, 1024], 8192), ("o", [8192], 8192), ("gate_up", [28672] * 2, 8192),
            ("down", [8192], 28672)],
}


def plan(parts, k, d=8, nbits=8, m=1, t=1):
    lib = _lib.load()
    arr, np_ = _lib.parts_array(parts)
    out = (ctypes.c_int * 9)()
    rc = lib.pq_b200_debug_lut_tma_plan(t, sum(parts), k, d, nbits, m, arr, np_, out)
    return None if rc else dict(zip(("cw", "cs", "S", "kg", "rows", "stages", "smem", "tiles", "grid"), out))


if __name__ == "__main__":
    for model, layers in SHAPES.items():
        for tp in (1, 2, 4, 8):
            for name, parts, k in layers:
                row = name in ("o", "down")
                p = [x // (1 if row else tp) for x in parts]
                if name == "qkv" and model == "70b" and tp > 1:
                    p = [8192 // tp, max(128, 1024 // tp), max(128, 1024 // tp)]
                print(model, f"tp{tp}", name, plan(p, k // tp if row else k))
