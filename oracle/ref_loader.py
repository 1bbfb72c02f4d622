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


_cached_layers = None


def load_reference_layers():
    """(RMSNorm, SiluAndMul): the reference's own layer classes out of layers/layernorm.py and
    layers/activation.py, executed unmodified under stubs, for their ``forward_native`` - the
    reference's CPU form of the ops the fused epilogue replaces.  The stubs:
    ``CustomOp`` is an nn.Module whose ``register`` decorator does nothing; ``current_platform``
    answers False to every ``is_*`` question (no native op is looked up); the TP helpers say 1 rank."""
    global _cached_layers
    if _cached_layers is not None:
        return _cached_layers
    if not available():
        raise FileNotFoundError(f"reference not present under {REF_ROOT}")
    saved = {k: v for k, v in sys.modules.items() if k == "vllm" or k.startswith("vllm.")}
    try:
        for name in ("vllm", "vllm.distributed", "vllm.model_executor", "vllm.model_executor.custom_op",
                     "vllm.model_executor.utils", "vllm.platforms", "vllm.utils", "vllm.model_executor.layers"):
            sys.modules[name] = types.ModuleType(name)

        class CustomOp(torch.nn.Module):              # model_executor/custom_op.py: dispatch only
            @classmethod
            def register(cls, name):
                return lambda op_cls: op_cls

            def forward(self, *args, **kwargs):
                return self.forward_native(*args, **kwargs)

        class _Platform:
            def __getattr__(self, name):
                if name.startswith("is_"):
                    return lambda *a, **k: False
                raise AttributeError(name)

        class LazyDict(dict):                          # vllm/utils.py: values are factories
            def __getitem__(self, key):
                return super().__getitem__(key)()

        sys.modules["vllm.model_executor.custom_op"].CustomOp = CustomOp
        dist = sys.modules["vllm.distributed"]
        dist.divide = lambda a, b: a // b
        dist.get_tensor_model_parallel_rank = lambda: 0
        dist.get_tensor_model_parallel_world_size = lambda: 1
        sys.modules["vllm.model_executor.utils"].set_weight_attrs = lambda w, attrs: None
        sys.modules["vllm.platforms"].current_platform = _Platform()
        sys.modules["vllm.utils"].LazyDict = LazyDict
        layers = "vllm/model_executor/layers"
        norm = _exec("vllm.model_executor.layers.layernorm", os.path.join(REF_ROOT, layers, "layernorm.py"))
        act = _exec("vllm.model_executor.layers.activation", os.path.join(REF_ROOT, layers, "activation.py"))
    finally:
        for k in [k for k in sys.modules if k == "vllm" or k.startswith("vllm.")]:
            del sys.modules[k]
        sys.modules.update(saved)
    _cached_layers = (norm.RMSNorm, act.SiluAndMul)
    return _cached_layers


_cached_linear = None


def load_reference_linear():
    """The reference's own TP linear layers with its own AQLM method: ``layers/linear.py``,
    ``parameter.py``, ``model_executor/utils.py``, ``quantization/base_config.py`` and
    ``quantization/aqlm.py`` executed unmodified under stubs.  Returns a namespace with
    ``linear``, ``aqlm``, ``vocab`` (layers/vocab_parallel_embedding.py) and ``mixtral``
    (models/mixtral_quant.py), ``llama`` (models/llama.py), ``registry`` (quantization/__init__.py) and
    ``base_config`` - the module
    objects - and ``set_tp(rank, world)``: the stubbed
    ``vllm.distributed`` answers with that simulated rank, so every rank's sharded ``create_weights`` and
    ``weight_loader`` can be run in ONE process.  The collectives are not stubbed with arithmetic: the
    all-gather raises, the all-reduce raises unless ``set_all_reduce(fn)`` installed a stand-in (the
    identity, to look at one simulated rank's partial result)."""
    global _cached_linear
    if _cached_linear is not None:
        return _cached_linear
    if not available():
        raise FileNotFoundError(f"reference not present under {REF_ROOT}")
    state = {"rank": 0, "world": 1}
    saved = {k: v for k, v in sys.modules.items() if k == "vllm" or k.startswith("vllm.")}
    try:
        for name in ("vllm", "vllm._custom_ops", "vllm.distributed", "vllm.logger", "vllm.platforms",
                     "vllm.model_executor", "vllm.model_executor.layers",
                     "vllm.model_executor.layers.quantization"):
            sys.modules[name] = types.ModuleType(name)
        sys.modules["vllm"]._custom_ops = sys.modules["vllm._custom_ops"]

        def divide(a, b):                               # distributed/utils.py: exact division or assert
            assert a % b == 0, f"{a} is not divisible by {b}"
            return a // b

        def no_collective(*a, **k):
            raise RuntimeError("reference collectives are not available under the stub")

        def split_tensor_along_last_dim(tensor, num_partitions, contiguous_split_chunks=False):
            return torch.split(tensor, divide(tensor.size(-1), num_partitions), dim=-1)

        dist = sys.modules["vllm.distributed"]
        dist.divide = divide
        dist.get_tensor_model_parallel_rank = lambda: state["rank"]
        dist.get_tensor_model_parallel_world_size = lambda: state["world"]
        dist.split_tensor_along_last_dim = split_tensor_along_last_dim
        def all_reduce(t):                              # set_all_reduce(fn) below; None: raise
            if state.get("all_reduce") is None:
                return no_collective()
            return state["all_reduce"](t)

        dist.tensor_model_parallel_all_gather = no_collective
        dist.tensor_model_parallel_all_reduce = all_reduce
        import logging
        sys.modules["vllm.logger"].init_logger = logging.getLogger

        class _Platform:
            simple_compile_backend = "eager"            # vocab_parallel_embedding.py:138 (torch.compile)

            def __getattr__(self, name):
                if name.startswith("is_"):
                    return lambda *a, **k: False
                raise AttributeError(name)

        sys.modules["vllm.platforms"].current_platform = _Platform()
        me = os.path.join(REF_ROOT, "vllm/model_executor")
        _exec("vllm.model_executor.utils", os.path.join(me, "utils.py"))
        sys_base = _exec("vllm.model_executor.layers.quantization.base_config",
                         os.path.join(REF_ROOT, _Q, "base_config.py"))
        _exec("vllm.model_executor.parameter", os.path.join(me, "parameter.py"))
        linear = _exec("vllm.model_executor.layers.linear", os.path.join(me, "layers/linear.py"))
        aqlm = _exec("vllm.model_executor.layers.quantization.aqlm", os.path.join(REF_ROOT, _Q, "aqlm.py"))
        registry = _exec("vllm.model_executor.layers.quantization", os.path.join(REF_ROOT, _Q, "__init__.py"))
        vocab = _exec("vllm.model_executor.layers.vocab_parallel_embedding",
                      os.path.join(me, "layers/vocab_parallel_embedding.py"))
        mixtral = _exec_mixtral_quant(me, dist, state)
        llama = _exec_llama(me)
    finally:
        for k in [k for k in sys.modules if k == "vllm" or k.startswith("vllm.")]:
            del sys.modules[k]
        sys.modules.update(saved)

    def set_tp(rank: int, world: int):
        state["rank"], state["world"] = rank, world

    def set_all_reduce(fn):
        """What the stubbed tensor_model_parallel_all_reduce does from now on (None: raise)."""
        state["all_reduce"] = fn

    _cached_linear = types.SimpleNamespace(linear=linear, aqlm=aqlm, vocab=vocab, mixtral=mixtral, llama=llama, registry=registry,
                                           base_config=sys_base, set_tp=set_tp,
                                           set_all_reduce=set_all_reduce)
    return _cached_linear


class _AnyModule(types.ModuleType):
    """A stub module whose every attribute is an empty placeholder class: enough for the
    ``from x import Name`` lines and type annotations of code paths that are never run."""

    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        cls = type(name, (), {})
        setattr(self, name, cls)
        return cls


def _exec_mixtral_quant(me: str, dist, state):
    """models/mixtral_quant.py (the reference's non-fused quantised Mixtral, the path an AQLM Mixtral
    takes) for its MixtralMLP / MixtralMoE classes.  Called inside load_reference_linear()'s stub
    environment; attention, rotary, sampler, embeddings ... are placeholders, never constructed."""
    dist.get_pp_group = lambda: None
    for name in ("vllm.attention", "vllm.config", "vllm.sequence", "vllm.model_executor.custom_op",
                 "vllm.model_executor.layers.logits_processor", "vllm.model_executor.layers.rotary_embedding",
                 "vllm.model_executor.layers.sampler",
                 "vllm.model_executor.model_loader", "vllm.model_executor.model_loader.weight_utils",
                 "vllm.model_executor.sampling_metadata", "vllm.model_executor.models",
                 "vllm.model_executor.models.interfaces", "vllm.model_executor.models.utils"):
        sys.modules[name] = _AnyModule(name)
    sys.modules["vllm.model_executor.models"].__path__ = []         # a package, for the relative imports

    class CustomOp(torch.nn.Module):
        @classmethod
        def register(cls, name):
            return lambda op_cls: op_cls

        def forward(self, *args, **kwargs):
            return self.forward_native(*args, **kwargs)

    sys.modules["vllm.model_executor.custom_op"].CustomOp = CustomOp
    base = sys.modules["vllm.model_executor.layers.quantization.base_config"]
    sys.modules["vllm.model_executor.layers.quantization"].QuantizationConfig = base.QuantizationConfig
    _exec("vllm.model_executor.layers.layernorm", os.path.join(me, "layers/layernorm.py"))
    return _exec("vllm.model_executor.models.mixtral_quant", os.path.join(me, "models/mixtral_quant.py"))


def _functions_of(path: str, names, namespace: dict):
    """Compile single top-level functions of a reference file (unmodified source, cut out by the AST) into
    ``namespace`` - for helper functions of modules whose other imports cannot be satisfied here."""
    import ast
    with open(path) as fh:
        tree = ast.parse(fh.read(), filename=path)
    body = [node for node in tree.body if isinstance(node, ast.FunctionDef) and node.name in names]
    assert {node.name for node in body} == set(names), (path, names)
    exec(compile(ast.Module(body=body, type_ignores=[]), path, "exec"), namespace)
    return [namespace[n] for n in names]


def _exec_llama(me: str):
    """models/llama.py for ``LlamaModel.load_weights`` (llama.py:369-433), the weight walk an AQLM Llama
    checkpoint goes through.  Called inside load_reference_linear()'s stub environment, after
    _exec_mixtral_quant() has put the placeholder modules in place.  ``default_weight_loader`` and
    ``maybe_remap_kv_scale_name`` are the reference's own functions, cut out of
    model_loader/weight_utils.py (that module's other imports - hub, safetensors, config - are not
    available under the stubs)."""
    import logging
    from typing import Optional
    comp = _AnyModule("vllm.compilation")
    dec = _AnyModule("vllm.compilation.decorators")
    dec.support_torch_compile = lambda cls: cls
    sys.modules["vllm.compilation"], sys.modules["vllm.compilation.decorators"] = comp, dec
    log = logging.getLogger("vllm.reference_stub")
    log.warning_once = log.warning
    ns = {"torch": torch, "Optional": Optional, "logger": log}
    loader, remap = _functions_of(os.path.join(me, "model_loader/weight_utils.py"),
                                  ["default_weight_loader", "maybe_remap_kv_scale_name"], ns)
    wu = sys.modules["vllm.model_executor.model_loader.weight_utils"]
    wu.default_weight_loader, wu.maybe_remap_kv_scale_name = loader, remap
    utils = sys.modules["vllm.model_executor.models.utils"]
    utils.is_pp_missing_parameter = lambda name, model: False          # no pipeline parallelism here
    class LazyDict(dict):                              # vllm/utils.py: values are factories
        def __getitem__(self, key):
            return super().__getitem__(key)()

    sys.modules["vllm.utils"] = _AnyModule("vllm.utils")
    sys.modules["vllm.utils"].LazyDict = LazyDict
    _exec("vllm.model_executor.layers.activation", os.path.join(me, "layers/activation.py"))
    return _exec("vllm.model_executor.models.llama", os.path.join(me, "models/llama.py"))
