"""ctypes binding of libpq_b200.so (C ABI in include/pq_b200.h).

There is no fallback: if the library is missing the import fails, and every
compute entry point needs a CUDA device.  Build it with
``make -C vllm_pq_b200/csrc`` or ``python -c 'import __graft_entry__ as g; g.build()'``.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import c_char_p, c_int, c_int64, c_void_p

_HERE = os.path.dirname(os.path.abspath(__file__))
# PQ_B200_LIB: another build of the same library (timing experiments); the default is the in-tree one
LIB_PATH = os.environ.get("PQ_B200_LIB") or os.path.join(_HERE, "libpq_b200.so")

PQ_F16, PQ_BF16 = 0, 1
ALGO_AUTO, ALGO_GEMV_DIRECT, ALGO_GEMV_LUT, ALGO_GEMM_TC, ALGO_GEMM_MMA, ALGO_GEMV_LUT_TMA = 0, 1, 2, 3, 4, 5
ALGO_NAMES = {0: "auto", 1: "gemv_direct", 2: "gemv_lut", 3: "gemm_tc", 4: "gemm_mma", 5: "gemv_lut_tma"}
ABI_VERSION = 2
MAX_PARTS = 8
AR_FLAG_BYTES = 4096
AR2_FLAG_BYTES = 8192
AR_MAX_WORLD = 8

# every symbol include/pq_b200.h declares: (restype, argtypes)
_I64P = ctypes.POINTER(c_int64)
_VPP = ctypes.POINTER(c_void_p)


class Epilogue(ctypes.Structure):
    """pq_b200_epilogue_t (include/pq_b200.h): extras of pq_b200_gemm_fused."""
    _fields_ = [("ar_world", ctypes.c_int32), ("ar_rank", ctypes.c_int32), ("ar_peer", _VPP),
                ("ar_state", c_void_p), ("ar_slot_elems", c_int64), ("residual", c_void_p),
                ("residual_stride", c_int64), ("ssq_out", c_void_p), ("ssq_slots", c_int64),
                ("norm_ssq", c_void_p), ("norm_weight", c_void_p), ("norm_slots", ctypes.c_int32),
                ("norm_eps", ctypes.c_float), ("input_silu_mul", ctypes.c_int32),
                ("codes_padded", c_void_p), ("codes_pitch", c_int64)]


SIGNATURES = {
    "pq_b200_abi_version": (c_int, []),
    "pq_b200_set_pdl": (None, [c_int]),
    "pq_b200_strerror": (c_char_p, [c_int]),
    "pq_b200_debug_set_trace": (None, [c_void_p]),
    "pq_b200_debug_last_launch": (None, [ctypes.POINTER(c_int)]),
    "pq_b200_debug_last_mma_plan": (None, [ctypes.POINTER(c_int)]),
    "pq_b200_debug_max_active_clusters": (c_int, [c_int, c_int]),
    "pq_b200_debug_lut_tma_slots": (c_int, [c_int, c_int, c_int]),
    "pq_b200_debug_lut_tma_plan": (c_int, [c_int64, c_int64, c_int64, c_int, c_int, c_int, _I64P, c_int,
                                           ctypes.POINTER(c_int)]),
    "pq_b200_dequant": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_int,
                                c_int, c_int, _I64P, c_int, c_int, c_void_p]),
    "pq_b200_workspace_bytes": (c_int64, [c_int, c_int64, c_int64, c_int64, c_int, c_int, c_int, c_int]),
    "pq_b200_gemm": (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                             c_int64, c_int64, c_int64, c_int64, c_int, c_int, c_int, _I64P, c_int,
                             c_int, c_int, c_void_p, c_int64, c_void_p]),
    "pq_b200_gemv_static": (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                    c_int64, c_int64, c_int64, c_int64, c_int, c_int, c_int, _I64P, c_int,
                                    c_int, c_void_p]),
    "pq_b200_gemm_silu_mul": (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                      c_int64, c_int64, c_int64, c_int64, c_int, c_int, c_int, _I64P, c_int,
                                      c_int, c_int, c_void_p, c_int64, c_void_p]),
    "pq_b200_gemm_silu_mul_workspace_bytes": (c_int64, [c_int, c_int64, c_int64, c_int64, c_int, c_int, c_int,
                                                        c_int]),
    "pq_b200_embedding": (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_int64,
                                  c_int, c_int, c_int, c_int, c_void_p]),
    "pq_b200_gemm_fused": (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                   c_int64, c_int64, c_int64, c_int64, c_int, c_int, c_int, _I64P, c_int,
                                   c_int, c_void_p, c_int64, c_void_p, ctypes.POINTER(Epilogue)]),
    "pq_b200_gemm_fused_supported": (c_int, [c_int64, c_int64, c_int64, c_int, c_int, c_int]),
    "pq_b200_gemm_fused_slots": (c_int64, [c_int64, c_int64, c_int64, c_int, c_int, c_int, _I64P, c_int, c_int64]),
    "pq_b200_fused_add_rms_norm": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_void_p, ctypes.c_float,
                                           c_int64, c_int64, c_int, c_void_p]),
    "pq_b200_silu_and_mul": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_int64, c_int64, c_int, c_void_p]),
    "pq_b200_algo_supported": (c_int, [c_int, c_int64, c_int64, c_int64, c_int, c_int, c_int]),
    "pq_b200_pick_algo": (c_int, [c_int64, c_int64, c_int64, c_int, c_int, c_int]),
    "pq_b200_allreduce_oneshot": (c_int, [c_void_p, c_void_p, c_int64, c_int, _VPP, _VPP, c_int, c_int,
                                          c_int64, c_void_p]),
    "pq_b200_allreduce_twostage": (c_int, [c_void_p, c_void_p, c_int64, c_int, _VPP, _VPP, _VPP, c_int, c_int,
                                           c_int64, c_void_p]),
    "pq_b200_malloc": (c_int, [c_int64, _VPP]),
    "pq_b200_free": (c_int, [c_void_p]),
    "pq_b200_ipc_get_handle": (c_int, [c_void_p, c_void_p]),
    "pq_b200_ipc_open_handle": (c_int, [c_void_p, _VPP]),
    "pq_b200_ipc_close_handle": (c_int, [c_void_p]),
    "pq_b200_memset_async": (c_int, [c_void_p, c_int, c_int64, c_void_p]),
}

_lib = None


def load() -> ctypes.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.isfile(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} not found: the CUDA library is required (no CPU fallback). "
            "Build it with `make -C vllm_pq_b200/csrc`.")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError if the ABI lost a symbol
        fn.restype, fn.argtypes = res, args
    got = lib.pq_b200_abi_version()
    if got != ABI_VERSION:
        raise ImportError(f"libpq_b200.so ABI {got}, python side expects {ABI_VERSION}")
    _lib = lib
    return lib


def strerror(code: int) -> str:
    return load().pq_b200_strerror(code).decode()


def check(code: int, what: str) -> None:
    """Non-zero C return -> RuntimeError, as TORCH_CHECK does in the reference
    (gemm_kernels.cu:539-540,594-595)."""
    if code != 0:
        raise RuntimeError(f"{what}: {strerror(code)} (code {code})")


def parts_array(partition_sizes):
    n = len(partition_sizes)
    if n < 1 or n > MAX_PARTS:
        raise RuntimeError(f"PQ supports 1..{MAX_PARTS} output partitions, got {n}")
    return (c_int64 * n)(*[int(p) for p in partition_sizes]), n
