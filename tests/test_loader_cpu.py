"""Checkpoint loader (vllm_pq_b200/loader.py; reference: llama.py:369-433 + safetensors iterator):
a synthetic AQLM-layout checkpoint on disk -> the fused / sharded PQ linears, on the CPU."""
import json
import os

import pytest
import torch
from safetensors.torch import save_file

import vllm_pq_b200 as P
from oracle import pq_oracle as O
from vllm_pq_b200 import loader as L
from vllm_pq_b200 import stack as S

SHAPE = S.LlamaShape("tiny", 128, 256, 4, 2, 2)          # hidden, intermediate, heads, kv heads, layers


def make_checkpoint(tmp_path, fmt, dtype=torch.float16, drop=None, extra_layer=False):
    b, d, m = fmt
    hd = SHAPE.head_dim
    mods = {"self_attn.q_proj": (SHAPE.heads * hd, SHAPE.hidden), "self_attn.k_proj": (SHAPE.kv_heads * hd, SHAPE.hidden),
            "self_attn.v_proj": (SHAPE.kv_heads * hd, SHAPE.hidden), "self_attn.o_proj": (SHAPE.hidden, SHAPE.heads * hd),
            "mlp.gate_proj": (SHAPE.intermediate, SHAPE.hidden), "mlp.up_proj": (SHAPE.intermediate, SHAPE.hidden),
            "mlp.down_proj": (SHAPE.hidden, SHAPE.intermediate)}
    full, tensors = {}, {}
    seed = 0
    for layer in range(SHAPE.layers + (1 if extra_layer else 0)):
        for mod, (n, k) in mods.items():
            seed += 1
            c = O.make_case([n], k, b, d, m, t=1, dtype=dtype, seed=seed)
            full[(layer, mod)] = c
            for key in ("codes", "codebooks", "scales"):
                tensors[f"model.layers.{layer}.{mod}.{key}"] = c[key].contiguous()
        tensors[f"model.layers.{layer}.input_layernorm.weight"] = torch.ones(SHAPE.hidden, dtype=dtype)
        tensors[f"model.layers.{layer}.self_attn.rotary_emb.inv_freq"] = torch.ones(hd // 2)
    tensors["model.embed_tokens.weight"] = torch.zeros(32, SHAPE.hidden, dtype=dtype)
    tensors["lm_head.weight"] = torch.zeros(32, SHAPE.hidden, dtype=dtype)
    if drop:
        tensors.pop(drop)
    names = sorted(tensors)
    half = len(names) // 2
    os.makedirs(tmp_path, exist_ok=True)
    save_file({n: tensors[n] for n in names[:half]}, os.path.join(tmp_path, "model-00001-of-00002.safetensors"))
    save_file({n: tensors[n] for n in names[half:]}, os.path.join(tmp_path, "model-00002-of-00002.safetensors"))
    with open(os.path.join(tmp_path, "config.json"), "w") as fh:
        json.dump({"model_type": "llama", "quantization_config": {
            "quant_method": "aqlm", "in_group_size": d, "nbits_per_codebook": b, "num_codebooks": m,
            "out_group_size": 1}}, fh)
    return full


@pytest.mark.parametrize("fmt", [(8, 8, 1), (8, 8, 2), (16, 8, 1)])
def test_load_checkpoint_fuses_and_places(tmp_path, fmt):
    b, d, m = fmt
    full = make_checkpoint(str(tmp_path), fmt)
    cfg = L.quant_config_from_hf(L.read_hf_config(str(tmp_path)))
    assert (cfg.in_group_size, cfg.nbits_per_codebook, cfg.num_codebooks, cfg.out_group_size) == (d, b, m, 1)
    model = S.PQLlamaLinearStack(SHAPE, cfg, torch.float16)
    loaded, skipped = L.load_checkpoint(model, str(tmp_path))
    assert len(loaded) == SHAPE.layers * 4 * 3                      # 4 linears x codes/codebooks/scales
    assert any("embed_tokens" in s for s in skipped) and any("lm_head" in s for s in skipped)
    assert any("input_layernorm" in s for s in skipped) and any("rotary_emb" in s for s in skipped)
    for li, layer in enumerate(model.layers):
        q, k, v = (full[(li, f"self_attn.{n}_proj")] for n in "qkv")
        assert torch.equal(layer.qkv_proj.codes, torch.cat([q["codes"], k["codes"], v["codes"]]))
        assert torch.equal(layer.qkv_proj.codebooks, torch.cat([q["codebooks"], k["codebooks"], v["codebooks"]]))
        assert torch.equal(layer.qkv_proj.scales, torch.cat([q["scales"], k["scales"], v["scales"]]))
        g, u = full[(li, "mlp.gate_proj")], full[(li, "mlp.up_proj")]
        assert torch.equal(layer.gate_up_proj.codes, torch.cat([g["codes"], u["codes"]]))
        assert torch.equal(layer.gate_up_proj.codebooks, torch.cat([g["codebooks"], u["codebooks"]]))
        for name, mod in (("o_proj", "self_attn.o_proj"), ("down_proj", "mlp.down_proj")):
            lin, c = getattr(layer, name), full[(li, mod)]
            for key in ("codes", "codebooks", "scales"):
                assert torch.equal(getattr(lin, key), c[key])
        # the fused layer dequantises (oracle) to the stack of its parts: partition sizes are right
        parts = list(layer.qkv_proj.codes.codebook_partition_sizes) if hasattr(layer.qkv_proj.codes, "codebook_partition_sizes") \
            else list(layer.qkv_proj.output_partition_sizes)
        w = O.dequantize_partitions(layer.qkv_proj.codes.data, layer.qkv_proj.codebooks.data, parts,
                                    layer.qkv_proj.scales.data)
        w_parts = torch.cat([O.dequantize_partitions(c["codes"], c["codebooks"], [c["codes"].shape[0]], c["scales"])
                             for c in (q, k, v)])
        assert torch.equal(w.view(torch.int16), w_parts.view(torch.int16))


def test_strict_reports_missing_and_unknown(tmp_path):
    make_checkpoint(str(tmp_path / "a"), (8, 8, 1), drop="model.layers.1.mlp.down_proj.scales")
    cfg = P.PQConfig(8, 8, 1, 1)
    model = S.PQLlamaLinearStack(SHAPE, cfg, torch.float16)
    with pytest.raises(KeyError, match="does not cover"):
        L.load_checkpoint(model, str(tmp_path / "a"))
    L.load_checkpoint(model, str(tmp_path / "a"), strict=False)
    make_checkpoint(str(tmp_path / "b"), (8, 8, 1), extra_layer=True)      # a layer the stack does not have
    with pytest.raises(KeyError):
        L.load_checkpoint(model, str(tmp_path / "b"))
    with pytest.raises(FileNotFoundError):
        list(L.safetensors_weights_iterator(str(tmp_path / "none")))


def test_quant_config_gate():
    with pytest.raises(ValueError):
        L.quant_config_from_hf({"model_type": "llama"})
    with pytest.raises(ValueError):
        L.quant_config_from_hf({"quantization_config": {"quant_method": "gptq"}})
