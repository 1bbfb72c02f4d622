"""The op layer's workspace bookkeeping (_custom_ops._workspace), host logic only:
one buffer per (device, stream), an outgrown buffer is RETIRED (kept alive - a CUDA graph captured
earlier holds its address), nothing is allocated under stream capture.  torch.cuda's stream queries are
stubbed; the buffers are CPU tensors."""
import types

import pytest
import torch

from vllm_pq_b200 import _custom_ops as ops


@pytest.fixture
def fake_cuda(monkeypatch):
    state = types.SimpleNamespace(stream=11, capturing=False)
    monkeypatch.setattr(torch.cuda, "current_stream", lambda device=None: types.SimpleNamespace(cuda_stream=state.stream))
    monkeypatch.setattr(torch.cuda, "current_device", lambda: 0)
    monkeypatch.setattr(torch.cuda, "is_current_stream_capturing", lambda: state.capturing)
    monkeypatch.setattr(ops, "_WORKSPACES", {})
    monkeypatch.setattr(ops, "_RETIRED", [])
    return state


def test_growth_retires_the_old_buffer(fake_cuda):
    dev = torch.device("cpu")
    a = ops._workspace(dev, 100)
    assert a.numel() == ops._WS_MIN_BYTES and a.dtype == torch.uint8 and not a.any()       # zeroed once
    assert ops._workspace(dev, ops._WS_MIN_BYTES) is a                                     # fits: same buffer
    ptr = a.data_ptr()
    b = ops._workspace(dev, ops._WS_MIN_BYTES + 1)
    assert b is not a and b.numel() == 2 * (ops._WS_MIN_BYTES + 1) and not b.any()
    assert ops._RETIRED == [a] and a.data_ptr() == ptr                                     # alive, address unchanged
    assert ops._workspace(dev, 100) is b


def test_one_buffer_per_stream(fake_cuda):
    dev = torch.device("cpu")
    a = ops._workspace(dev, 100)
    fake_cuda.stream = 12
    b = ops._workspace(dev, 100)
    assert b is not a and a.data_ptr() != b.data_ptr()
    fake_cuda.stream = 11
    assert ops._workspace(dev, 100) is a and ops._RETIRED == []


def test_nothing_is_allocated_under_capture(fake_cuda):
    dev = torch.device("cpu")
    a = ops._workspace(dev, 100)
    fake_cuda.capturing = True
    assert ops._workspace(dev, 200) is a                       # the buffer from the eager warm-up serves the capture
    assert ops._workspace(dev, a.numel() + 1) is None          # too small: no workspace for this launch, a stays
    fake_cuda.stream = 13
    assert ops._workspace(dev, 100) is None                    # a stream that never warmed up
    assert ops._RETIRED == [] and list(ops._WORKSPACES.values()) == [a]
