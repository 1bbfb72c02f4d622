"""Entry point for vLLM's ``vllm.general_plugins`` group (reference
vllm/plugins/__init__.py:50-82): registers the PQ method inside a running vLLM.

    [project.entry-points."vllm.general_plugins"]
    pq_b200 = "vllm_pq_b200.vllm_plugin:register"

Stand-alone (no vLLM) the package uses its own restatement of the plugin ABCs
(``plugin_api.py``).  Inside vLLM the config and the linear method must be
subclasses of vLLM's own ``QuantizationConfig`` / ``LinearMethodBase``
(``register_quantization_config`` checks ``issubclass``,
quantization/__init__.py:62-67) and ``get_quant_method`` must recognise vLLM's
``LinearBase`` (aqlm.py:213-217).  ``register()`` asks ``pq.make_classes`` for the same
method bodies (``PQConfigBody`` / ``PQLinearMethodBody``) mounted on vLLM's bases; nothing else
differs.  Tested against the vLLM installed in the image: tests/test_vllm_plugin_cpu.py.
"""
from __future__ import annotations

_REGISTERED = None


def register(name: str = "pq"):
    """Idempotent; returns the registered config class (a subclass of vLLM's QuantizationConfig whose
    ``get_quant_method`` hands vLLM's LinearBase layers a subclass of vLLM's LinearMethodBase)."""
    global _REGISTERED
    if _REGISTERED is not None:
        return _REGISTERED
    from vllm.model_executor.layers.linear import LinearBase, LinearMethodBase
    from vllm.model_executor.layers.quantization import register_quantization_config
    from vllm.model_executor.layers.quantization.base_config import QuantizationConfig

    from .pq import make_classes

    from vllm.model_executor.layers.vocab_parallel_embedding import VocabParallelEmbedding

    config_cls, _ = make_classes(QuantizationConfig, LinearMethodBase,
                                 lambda layer: isinstance(layer, LinearBase), name,
                                 is_vocab=lambda layer: isinstance(layer, VocabParallelEmbedding))
    _REGISTERED = register_quantization_config(name)(config_cls)
    return _REGISTERED
