"""Tensor-parallel state and the reduction hook of the row-parallel PQ linears.

One process per GPU; ``torch.distributed`` (NCCL on GPUs, gloo in CPU tests)
is the plumbing.  Mirrors the one call chain of the reference that the PQ path
reaches:
  tensor_model_parallel_all_reduce   vllm/distributed/communication_op.py:11-13
  GroupCoordinator.all_reduce        vllm/distributed/parallel_state.py:287-313
  CudaCommunicator.all_reduce        device_communicators/cuda_communicator.py:52-72
  CustomAllreduce                    device_communicators/custom_all_reduce.py:55-298
Decode-sized messages go through the one-shot NVLink kernel in
csrc/pq_allreduce.cu (each rank pushes its vector into every peer's buffer, one
flag barrier, local fp32 sum in rank order) - or, at 1..8 tokens, ride in the
row-parallel GEMV's own epilogue (csrc/pq_gemv_lut2.cu); messages up to 64 MiB
(prefill) through the two-stage reduce-scatter + all-gather kernel over the same
peer mappings, which is CUDA-graph capturable; anything larger through NCCL.
"""
from __future__ import annotations

import ctypes
import os
from typing import List, Optional

import torch
import torch.distributed as dist

_TP_GROUP: Optional[dist.ProcessGroup] = None
_TP_RANK = 0
_TP_WORLD = 1
_ONESHOT: Optional["OneShotAllReduce"] = None


def init_tensor_parallel(group: Optional[dist.ProcessGroup] = None, oneshot: Optional[bool] = None,
                         max_oneshot_bytes: int = 1 << 20) -> None:
    """All ranks of ``group`` (default: the world) form one TP group
    (parallel_state.py:865-923 with pp = dp = 1)."""
    global _TP_GROUP, _TP_RANK, _TP_WORLD, _ONESHOT
    if not dist.is_initialized():
        _TP_GROUP, _TP_RANK, _TP_WORLD, _ONESHOT = None, 0, 1, None
        return
    _TP_GROUP = group if group is not None else dist.group.WORLD
    _TP_RANK = dist.get_rank(_TP_GROUP)
    _TP_WORLD = dist.get_world_size(_TP_GROUP)
    _ONESHOT = None
    if oneshot is None:
        oneshot = os.environ.get("PQ_ONESHOT_ALLREDUCE", "1") != "0"
    if (oneshot and _TP_WORLD in (2, 4, 8) and torch.cuda.is_available()
            and dist.get_backend(_TP_GROUP) == "nccl"):
        _ONESHOT = _setup_oneshot(_TP_GROUP, max_oneshot_bytes)


def _setup_oneshot(group: dist.ProcessGroup, max_bytes: int) -> Optional["OneShotAllReduce"]:
    """Peer-memory all-reduce for the whole group or for nobody: a rank that went to NCCL while its
    peers spin in the NVLink kernel would hang the group, so every step that can fail on one rank
    only (same node? peer access? cudaIpcOpenMemHandle?) ends in an agreed MIN over the group.
    Checks follow custom_all_reduce.py:86-147 (single node, full P2P)."""
    import socket
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    where = [None] * world
    dist.all_gather_object(where, (socket.gethostname(), torch.cuda.current_device()), group=group)
    ok, why = True, ""
    if len({h for h, _ in where}) != 1:
        ok, why = False, "ranks span more than one node"
    else:
        me = torch.cuda.current_device()
        for r, (_, d) in enumerate(where):
            if r != rank and (d == me or not torch.cuda.can_device_access_peer(me, d)):
                ok, why = False, f"no peer access from device {me} to device {d}"
                break
    comm = None
    if _agree(ok, group):
        try:
            comm = OneShotAllReduce(group, max_bytes)
        except Exception as exc:                       # noqa: BLE001 - reported below, agreed on by all ranks
            ok, why = False, str(exc)
        if not _agree(ok and comm is not None, group):
            if comm is not None:
                comm.close()
            comm = None
    if comm is None and rank == 0:
        print(f"[vllm_pq_b200] one-shot all-reduce disabled on every rank ({why or 'a peer could not set it up'}); using NCCL")
    return comm


def _agree(ok: bool, group: dist.ProcessGroup) -> bool:
    device = "cuda" if dist.get_backend(group) == "nccl" else "cpu"       # gloo: the CPU tests
    flag = torch.tensor([1 if ok else 0], device=device, dtype=torch.int32)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
    return bool(flag.item())


def destroy_tensor_parallel() -> None:
    global _TP_GROUP, _TP_RANK, _TP_WORLD, _ONESHOT
    if _ONESHOT is not None:
        _ONESHOT.close()
    _TP_GROUP, _TP_RANK, _TP_WORLD, _ONESHOT = None, 0, 1, None


def get_tensor_model_parallel_world_size() -> int:
    return _TP_WORLD


def get_tensor_model_parallel_rank() -> int:
    return _TP_RANK


def oneshot_enabled() -> bool:
    return _ONESHOT is not None


def fused_allreduce_comm() -> Optional["OneShotAllReduce"]:
    """The peer-memory communicator the fused GEMV + all-reduce epilogue uses, or None (TP = 1, no
    peer access, or PQ_FUSED_ALLREDUCE=0)."""
    if _ONESHOT is None or os.environ.get("PQ_FUSED_ALLREDUCE", "1") == "0":
        return None
    return _ONESHOT


def tensor_model_parallel_all_reduce(input_: torch.Tensor) -> torch.Tensor:
    """All-reduce SUM across the TP group; passthrough at world size 1
    (parallel_state.py:303-304)."""
    if _TP_WORLD == 1:
        return input_
    if _ONESHOT is not None and _ONESHOT.eligible(input_):
        return _ONESHOT.all_reduce(input_)
    if (_ONESHOT is not None and _ONESHOT.eligible_twostage(input_)
            and os.environ.get("PQ_TWOSTAGE_ALLREDUCE", "1") != "0"):
        return _ONESHOT.all_reduce_twostage(input_)      # above the one-shot limit: still one capturable kernel
    dist.all_reduce(input_, group=_TP_GROUP)     # input_ is a fresh GEMM output: in place is safe
    return input_


def tensor_model_parallel_all_gather(input_: torch.Tensor, dim: int = -1) -> torch.Tensor:
    """Only reached with gather_output=True, which the Llama layers never set
    (linear.py:426-428,787)."""
    if _TP_WORLD == 1:
        return input_
    dim = dim % input_.dim()
    parts = [torch.empty_like(input_) for _ in range(_TP_WORLD)]
    dist.all_gather(parts, input_.contiguous(), group=_TP_GROUP)
    return torch.cat(parts, dim=dim)


class OneShotAllReduce:
    """Host side of pq_b200_allreduce_oneshot: owns this rank's staging + flag
    buffers (cudaMalloc'd by the C library so the IPC handle maps the exact base),
    exchanges cudaIpc handles over the process group, keeps the epoch counter.
    Eligibility follows custom_all_reduce.py:247-260: fp16/bf16, 16-byte multiple,
    at most ``max_bytes``; world size 2/4/8 on one NVSwitch node."""

    # fused GEMV + all-reduce (csrc/pq_gemv_lut2.cu): receive buffer of {fp32, epoch} packets,
    # [2 parities][world][LL_SLOT_ELEMS]; 8 tokens (the table GEMV's limit) x 16384 outputs
    LL_SLOT_ELEMS = 8 * 16384
    # two-stage all-reduce (csrc/pq_allreduce.cu) for messages above `max_bytes`: 2048 tokens x 8192
    # hidden x 2 bytes = 32 MiB is Llama-2-70B's prefill partial sum; beyond 64 MiB NCCL takes over
    TWOSTAGE_MAX_BYTES = 64 << 20

    def __init__(self, group: dist.ProcessGroup, max_bytes: int):
        from . import _lib
        self._libmod = _lib
        self.lib = _lib.load()
        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self.max_bytes = int(max_bytes)
        self._opened: List[int] = []
        # two staging slots (alternating by the device-side epoch parity) so that a fast
        # rank's next push cannot overwrite data a slow rank is still summing
        self.data = self.flags = self.ll_data = self.ll_state = None
        self.stage2 = self.res2 = self.flags2 = None
        self.res2_bytes = 2 * ((self.TWOSTAGE_MAX_BYTES // self.world + 255) // 256 * 256)
        mine = None
        try:
            self.data = self._malloc(2 * self.world * self.max_bytes)
            self.flags = self._malloc(_lib.AR_FLAG_BYTES)
            ll_bytes = 2 * self.world * self.LL_SLOT_ELEMS * 8
            self.ll_data = self._malloc(ll_bytes)
            self.ll_state = self._malloc(256)        # {epoch, ticket}: local only, never mapped by peers
            self.stage2 = self._malloc(self.TWOSTAGE_MAX_BYTES)
            self.res2 = self._malloc(self.res2_bytes)
            self.flags2 = self._malloc(_lib.AR2_FLAG_BYTES)
            stream = torch.cuda.current_stream().cuda_stream
            _lib.check(self.lib.pq_b200_memset_async(self.flags, 0, _lib.AR_FLAG_BYTES, stream), "memset")
            _lib.check(self.lib.pq_b200_memset_async(self.ll_data, 0, ll_bytes, stream), "memset")
            _lib.check(self.lib.pq_b200_memset_async(self.ll_state, 0, 256, stream), "memset")
            _lib.check(self.lib.pq_b200_memset_async(self.flags2, 0, _lib.AR2_FLAG_BYTES, stream), "memset")
            torch.cuda.synchronize()
            mine = (self._handle(self.data), self._handle(self.flags), self._handle(self.ll_data),
                    self._handle(self.stage2), self._handle(self.res2), self._handle(self.flags2))
        except Exception as exc:                       # noqa: BLE001 - the collective below must still run
            mine = None
            local_error = exc
        handles = [None] * self.world
        dist.all_gather_object(handles, mine, group=group)
        if mine is None:
            raise local_error
        if any(h is None for h in handles):
            raise RuntimeError("a peer could not allocate its staging buffers")
        self.peer_data = (ctypes.c_void_p * self.world)()
        self.peer_flags = (ctypes.c_void_p * self.world)()
        self.peer_ll = (ctypes.c_void_p * self.world)()
        self.peer_stage2 = (ctypes.c_void_p * self.world)()
        self.peer_res2 = (ctypes.c_void_p * self.world)()
        self.peer_flags2 = (ctypes.c_void_p * self.world)()
        for r, (hd, hf, hl, hs, hr, hf2) in enumerate(handles):
            if r == self.rank:
                self.peer_data[r], self.peer_flags[r], self.peer_ll[r] = self.data, self.flags, self.ll_data
                self.peer_stage2[r], self.peer_res2[r], self.peer_flags2[r] = self.stage2, self.res2, self.flags2
            else:
                self.peer_data[r] = self._open(hd)
                self.peer_flags[r] = self._open(hf)
                self.peer_ll[r] = self._open(hl)
                self.peer_stage2[r] = self._open(hs)
                self.peer_res2[r] = self._open(hr)
                self.peer_flags2[r] = self._open(hf2)
        # no barrier here: the caller (_setup_oneshot) ends with a group-wide agreement, which also
        # orders every rank's memset of its flags before any peer's first kernel

    def _malloc(self, nbytes: int) -> int:
        p = ctypes.c_void_p()
        self._libmod.check(self.lib.pq_b200_malloc(nbytes, ctypes.byref(p)), "pq_b200_malloc")
        return p.value

    def _handle(self, ptr: int) -> bytes:
        buf = ctypes.create_string_buffer(64)
        self._libmod.check(self.lib.pq_b200_ipc_get_handle(ptr, buf), "ipc_get_handle")
        return buf.raw

    def _open(self, handle: bytes) -> int:
        p = ctypes.c_void_p()
        buf = ctypes.create_string_buffer(handle, 64)
        self._libmod.check(self.lib.pq_b200_ipc_open_handle(buf, ctypes.byref(p)), "ipc_open_handle")
        self._opened.append(p.value)
        return p.value

    def fill_epilogue(self, ep) -> None:
        """all-reduce fields of a _lib.Epilogue (pq_b200_epilogue_t)."""
        ep.ar_world, ep.ar_rank = self.world, self.rank
        ep.ar_peer = ctypes.cast(self.peer_ll, ctypes.POINTER(ctypes.c_void_p))
        ep.ar_state = self.ll_state
        ep.ar_slot_elems = self.LL_SLOT_ELEMS

    def eligible(self, t: torch.Tensor) -> bool:
        nbytes = t.numel() * t.element_size()
        return (t.is_cuda and t.dtype in (torch.float16, torch.bfloat16) and t.is_contiguous()
                and nbytes % 16 == 0 and nbytes <= self.max_bytes and t.data_ptr() % 16 == 0)

    def eligible_twostage(self, t: torch.Tensor) -> bool:
        nbytes = t.numel() * t.element_size()
        return (t.is_cuda and t.dtype in (torch.float16, torch.bfloat16) and t.is_contiguous()
                and nbytes % 16 == 0 and nbytes <= self.TWOSTAGE_MAX_BYTES and t.data_ptr() % 16 == 0)

    def all_reduce_twostage(self, t: torch.Tensor) -> torch.Tensor:
        out = torch.empty_like(t)
        dt = self._libmod.PQ_F16 if t.dtype == torch.float16 else self._libmod.PQ_BF16
        rc = self.lib.pq_b200_allreduce_twostage(t.data_ptr(), out.data_ptr(), t.numel(), dt, self.peer_stage2,
                                                 self.peer_res2, self.peer_flags2, self.rank, self.world,
                                                 self.TWOSTAGE_MAX_BYTES, torch.cuda.current_stream().cuda_stream)
        self._libmod.check(rc, "pq_b200_allreduce_twostage")
        return out

    def all_reduce(self, t: torch.Tensor) -> torch.Tensor:
        out = torch.empty_like(t)
        dt = self._libmod.PQ_F16 if t.dtype == torch.float16 else self._libmod.PQ_BF16
        rc = self.lib.pq_b200_allreduce_oneshot(t.data_ptr(), out.data_ptr(), t.numel(), dt,
                                                self.peer_data, self.peer_flags, self.rank, self.world,
                                                self.max_bytes,
                                                torch.cuda.current_stream().cuda_stream)
        self._libmod.check(rc, "pq_b200_allreduce_oneshot")
        return out

    def close(self) -> None:
        try:
            torch.cuda.synchronize()
            for p in self._opened:
                self.lib.pq_b200_ipc_close_handle(p)
            for p in (self.data, self.flags, self.ll_data, self.ll_state, self.stage2, self.res2, self.flags2):
                if p:
                    self.lib.pq_b200_free(p)
        except Exception:
            pass
        self._opened = []
