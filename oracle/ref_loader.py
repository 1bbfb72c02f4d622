"""Load the reference's own ``aqlm.py`` on the CPU.  TEST INFRASTRUCTURE ONLY.

``import vllm`` from /root/reference does not work in this image (SURVEY.md
section 8c), but ``aqlm.py`` needs only torch plus four vllm imports
(aqlm.py:13-17).  Those are stubbed here; the real ``base_config.py`` is
exec-loaded under its dotted name.  Nothing is copied: both files are read
from where they lie.  /root/reference exists only in the build container, so
``available()`` is False on the GPU box and every caller must cope (the golden
fixtures under tests/golden/ are what travels).
"""
from __future__ import annotations

import importlib.util
import os
import sys
import types

import torch

REF_ROOT = os.environ.get("PQ_REFERENCE_ROOT", "/root/reference")
_Q = "vllm/model_executor/layers/quantization"


def available() -> bool:
    return os.path.isfile(os.path.join(REF_ROOT, _Q, "aqlm.py"))


def _exec(dotted: str, path: str):
    spec = importlib.util.spec_from_file_location(dotted, path)
    mod = importlib.util.module_from_spec(spec)
    sys.modules[dotted] = mod
    spec.loader.exec_module(mod)
    return mod


_cached = None


def load_reference_aqlm():
    """Returns the reference module object (aqlm.py executed unmodified)."""
    global _cached
    if _cached is not None:
        return _cached
    if not available():
        raise FileNotFoundError(f"reference not present under {REF_ROOT}")
    saved = {k: v for k, v in sys.modules.items() if k == "vllm" or k.startswith("vllm.")}
    try:
        for name in ("vllm", "vllm._custom_ops", "vllm.model_executor",
                     "vllm.model_executor.layers",
                     "vllm.model_executor.layers.quantization",
                     "vllm.model_executor.layers.linear",
                     "vllm.model_executor.utils"):
            sys.modules[name] = types.ModuleType(name)
        sys.modules["vllm"]._custom_ops = sys.modules["vllm._custom_ops"]

        lin = sys.modules["vllm.model_executor.layers.linear"]
        base = _exec("vllm.model_executor.layers.quantization.base_config",
                     os.path.join(REF_ROOT, _Q, "base_config.py"))

        class LinearBase(torch.nn.Module):
            pass

        class LinearMethodBase(base.QuantizeMethodBase):
            pass

        lin.LinearBase, lin.LinearMethodBase = LinearBase, LinearMethodBase

        def set_weight_attrs(weight, attrs):      # model_executor/utils.py:13-44
            if attrs is None:
                return
            for key, value in attrs.items():
                assert not hasattr(weight, key), key
                setattr(weight, key, value)

        sys.modules["vllm.model_executor.utils"].set_weight_attrs = set_weight_attrs
        mod = _exec("vllm.model_executor.layers.quantization.aqlm",
                    os.path.join(REF_ROOT, _Q, "aqlm.py"))
        mod.LinearBase = LinearBase
    finally:
        for k in [k for k in sys.modules if k == "vllm" or k.startswith("vllm.")]:
            del sys.modules[k]
        sys.modules.update(saved)
    _cached = mod
    return mod
