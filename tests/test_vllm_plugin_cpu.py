"""The in-vLLM side of the boundary, against the vLLM installed in the image (CPU only).

Mirrors the reference's own plugin test (tests/quantization/test_register_quantization_config.py:49-97):
register -> look the class up by name -> from_config -> get_quant_method on a vLLM linear layer ->
create_weights, plus the duplicate-name rule of the reference registry (quantization/__init__.py:50-54,
restated in plugin_api.register_quantization_config).
"""
import pytest
import torch

import vllm_pq_b200 as P

HF = {"in_group_size": 8, "nbits_per_codebook": 8, "num_codebooks": 2, "out_group_size": 1}


@pytest.fixture(scope="module")
def vllm_cls():
    pytest.importorskip("vllm")
    from vllm_pq_b200 import vllm_plugin
    return vllm_plugin.register()


def test_register_is_idempotent_and_visible(vllm_cls):
    from vllm.model_executor.layers.quantization import QUANTIZATION_METHODS, get_quantization_config
    from vllm.model_executor.layers.quantization.base_config import QuantizationConfig
    from vllm_pq_b200 import vllm_plugin
    assert vllm_plugin.register() is vllm_cls
    assert "pq" in QUANTIZATION_METHODS
    assert get_quantization_config("pq") is vllm_cls
    assert issubclass(vllm_cls, QuantizationConfig)
    assert vllm_cls is not P.PQConfig                       # stand-alone class stays on our own ABCs


def test_from_config_constructs(vllm_cls):
    cfg = vllm_cls.from_config(dict(HF))                    # r1: TypeError from a re-based zero-arg super()
    assert (cfg.in_group_size, cfg.nbits_per_codebook, cfg.num_codebooks, cfg.out_group_size) == (8, 8, 2, 1)
    assert cfg.pack_factor == 8 and cfg.get_name() == "pq"
    assert cfg.packed_modules_mapping == {}                 # vLLM's QuantizationConfig.__init__ ran
    assert vllm_cls.get_min_capability() == 100
    assert vllm_cls.get_supported_act_dtypes() == [torch.half, torch.bfloat16]
    with pytest.raises(ValueError):
        vllm_cls.from_config({"in_group_size": 8})
    assert vllm_cls.override_quantization_method({"quant_method": "aqlm"}, "pq") == "pq"
    assert vllm_cls.override_quantization_method({"quant_method": "aqlm"}, None) is None


def test_quant_method_on_vllm_linear(vllm_cls):
    from vllm.model_executor.layers.linear import LinearMethodBase, ReplicatedLinear
    cfg = vllm_cls.from_config(dict(HF))
    layer = ReplicatedLinear(64, 32, bias=False, params_dtype=torch.float16, quant_config=cfg,
                             prefix="model.layers.0.mlp.down_proj", disable_tp=True)
    qm = layer.quant_method
    assert isinstance(qm, LinearMethodBase) and type(qm).__name__ == "PQLinearMethod"
    shapes = {n: (tuple(p.shape), p.dtype) for n, p in layer.named_parameters()}
    assert shapes == {"codes": ((32, 8, 2), torch.int8), "codebooks": ((2, 256, 1, 8), torch.float16),
                      "scales": ((32, 1, 1, 1), torch.float16)}       # aqlm.py:253-316
    assert layer.codes.pack_factor == 8 and layer.codes.packed_dim == 1
    assert layer.codebooks.is_metadata and layer.codebooks.output_partition_sizes == [32]
    # vLLM's own (unmodified) weight loader fills the parameters
    w = torch.randint(-128, 128, (32, 8, 2), dtype=torch.int8)
    layer.weight_loader(layer.codes, w)
    assert torch.equal(layer.codes.data, w)
    # embeddings / anything that is not a LinearBase stays unquantized (aqlm.py:213-217)
    assert cfg.get_quant_method(torch.nn.Embedding(4, 4), "embed") is None
    with pytest.raises(ValueError):
        qm.create_weights(torch.nn.Module(), 60, [32], 60, 32, torch.float16)   # K % d != 0
    with pytest.raises(ValueError):
        qm.create_weights(torch.nn.Module(), 64, [32], 64, 32, torch.float32)


def test_apply_has_no_cpu_path(vllm_cls):
    from vllm.model_executor.layers.linear import ReplicatedLinear
    cfg = vllm_cls.from_config(dict(HF))
    layer = ReplicatedLinear(64, 32, bias=False, params_dtype=torch.float16, quant_config=cfg, prefix="x",
                             disable_tp=True)
    with pytest.raises((RuntimeError, NotImplementedError)):
        layer.quant_method.apply(layer, torch.zeros(1, 64, dtype=torch.float16))


def test_own_registry_rejects_duplicates_and_non_configs():
    with pytest.raises(ValueError):                          # "pq" is taken by the package itself
        P.register_quantization_config("pq")(P.PQConfig)
    with pytest.raises(ValueError):
        P.register_quantization_config("pq_other")(object)
    assert P.get_quantization_config("pq") is P.PQConfig
    with pytest.raises(ValueError):
        P.get_quantization_config("nope")


def test_make_classes_mounts_the_same_bodies():
    from vllm_pq_b200 import pq
    from vllm_pq_b200.plugin_api import LinearMethodBase, QuantizationConfig
    cfg_cls, m_cls = pq.make_classes(QuantizationConfig, LinearMethodBase, lambda l: True, "pq_x")
    cfg = cfg_cls(8, 8, 1, 1)
    assert cfg.get_name() == "pq_x" and isinstance(cfg.get_quant_method(object(), ""), m_cls)
    assert m_cls.apply is pq.PQLinearMethodBody.apply and cfg_cls.__init__ is pq.PQConfigBody.__init__
