"""Entry point for vLLM's ``vllm.general_plugins`` group (reference
vllm/plugins/__init__.py:50-82): registers the PQ method inside a running vLLM.

    [project.entry-points."vllm.general_plugins"]
    pq_b200 = "vllm_pq_b200.vllm_plugin:register"

Stand-alone (no vLLM) the package uses its own restatement of the plugin ABCs
(``plugin_api.py``).  Inside vLLM the config and the linear method must be
subclasses of vLLM's own ``QuantizationConfig`` / ``LinearMethodBase``
(``register_quantization_config`` checks ``issubclass``,
quantization/__init__.py:62-67) and ``get_quant_method`` must recognise vLLM's
``LinearBase`` (aqlm.py:213-217).  ``register()`` builds those subclasses from
the same method bodies; nothing else differs.
"""
from __future__ import annotations

_REGISTERED = None


def register(name: str = "pq"):
    """Idempotent; returns the registered config class."""
    global _REGISTERED
    if _REGISTERED is not None:
        return _REGISTERED
    from vllm.model_executor.layers.linear import LinearBase, LinearMethodBase
    from vllm.model_executor.layers.quantization import register_quantization_config
    from vllm.model_executor.layers.quantization.base_config import QuantizationConfig

    from .pq import PQConfig, PQLinearMethod

    skip = ("__dict__", "__weakref__", "__module__", "__doc__", "__abstractmethods__", "_abc_impl")

    def rebase(cls, base, new_name):
        body = {k: v for k, v in vars(cls).items() if k not in skip}
        return type(new_name, (base,), body)

    method_cls = rebase(PQLinearMethod, LinearMethodBase, "PQLinearMethod")

    def get_quant_method(self, layer, prefix):
        return method_cls(self) if isinstance(layer, LinearBase) else None

    config_cls = rebase(PQConfig, QuantizationConfig, "PQConfig")
    config_cls.get_quant_method = get_quant_method
    _REGISTERED = register_quantization_config(name)(config_cls)
    return _REGISTERED
