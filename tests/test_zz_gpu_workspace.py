"""Capture, then growth, then replay: a CUDA graph captured while the stream's
workspace was small must stay valid after the op layer has replaced that workspace by a larger one.
Written after the round's GPU minutes were spent - hence near the end of the collection; the host logic
is covered on the CPU in tests/test_workspace_cpu.py."""
import pytest
import torch

import vllm_pq_b200 as P
from oracle import pq_oracle as O
from tests.test_gpu_parity import check_output, dev
from vllm_pq_b200 import _custom_ops as ops
from vllm_pq_b200 import _lib

pytestmark = pytest.mark.gpu


def test_graph_survives_workspace_growth():
    # down_proj of Llama-2-7B at one token: the table GEMV splits K over CTA groups, whose partial sums and
    # ticket counters go through the workspace
    c = O.make_case([4096], 11008, 8, 8, 1, t=1, dtype=torch.float16, seed=91)
    g = dev(c)
    args = (g["x"], g["codes"], g["codebooks"], g["scales"], [4096], None)
    assert _lib.load().pq_b200_workspace_bytes(_lib.ALGO_AUTO, 1, 4096, 11008, 8, 8, 1, 1) > 0
    eager = P.ops.pq_gemm(*args)
    check_output(eager, c, "eager")
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        P.ops.pq_gemm(*args)                                   # warm-up on the capture stream: its workspace
        old = ops._workspace(g["x"].device, 1)
        assert old is not None
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=s):
            out = P.ops.pq_gemm(*args)
        graph.replay()
        torch.cuda.synchronize()
        assert torch.equal(out, eager)
        # growth on the same stream: the old buffer is retired, not freed
        new = ops._workspace(g["x"].device, old.numel() + 1)
        assert new is not old and new.numel() > old.numel() and any(r is old for r in ops._RETIRED)
        # memory the allocator would hand out again if the old buffer had been dropped
        junk = [torch.full((old.numel(),), 255, dtype=torch.uint8, device="cuda") for _ in range(3)]
        after = P.ops.pq_gemm(*args)                           # eager, through the new buffer
        out.zero_()
        graph.replay()                                         # captured with the old buffer's address
        torch.cuda.synchronize()
    assert torch.equal(after, eager) and torch.equal(out, eager)
    assert not old[:65536].any() and not new[:65536].any()     # every launch leaves the ticket area at zero
    del junk
