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


# ---- vocabulary layers and Mixtral experts -----------------------------------------
class TinyMoEModel(torch.nn.Module):
    """embed_tokens + lm_head + two layers of {attention projections, MoE block}: the module names an HF
    Mixtral checkpoint uses (mixtral_quant.py:300-330 in the reference)."""

    def __init__(self, cfg, vocab=100, hidden=128, ffn=256, experts=4, tie=False):
        super().__init__()
        from vllm_pq_b200.embedding import ParallelLMHead, VocabParallelEmbedding
        from vllm_pq_b200.moe import PQMixtralMoE
        kw = dict(params_dtype=torch.float16, quant_config=cfg)
        self.embed_tokens = VocabParallelEmbedding(vocab, hidden, prefix="model.embed_tokens", **kw)
        self.lm_head = ParallelLMHead(vocab, hidden, prefix="lm_head", **kw)
        if tie:
            self.lm_head.tie_weights(self.embed_tokens)
        self.layers = torch.nn.ModuleList()
        for _ in range(2):
            layer = torch.nn.Module()
            layer.block_sparse_moe = PQMixtralMoE(experts, 2, hidden, ffn, **kw)
            self.layers.append(layer)


def moe_checkpoint(fmt, vocab=100, hidden=128, ffn=256, experts=4, lm_head=True):
    b, d, m = fmt
    tensors, full, seed = {}, {}, 100

    def put(prefix, n, k):
        nonlocal seed
        seed += 1
        c = O.make_case([n], k, b, d, m, t=1, dtype=torch.float16, seed=seed)
        full[prefix] = c
        for key in ("codes", "codebooks", "scales"):
            tensors[f"{prefix}.{key}"] = c[key].contiguous()

    put("model.embed_tokens", vocab, hidden)
    if lm_head:
        put("lm_head", vocab, hidden)
    for layer in range(2):
        for e in range(experts):
            put(f"model.layers.{layer}.block_sparse_moe.experts.{e}.w1", ffn, hidden)
            put(f"model.layers.{layer}.block_sparse_moe.experts.{e}.w2", hidden, ffn)
            put(f"model.layers.{layer}.block_sparse_moe.experts.{e}.w3", ffn, hidden)
        tensors[f"model.layers.{layer}.block_sparse_moe.gate.weight"] = torch.randn(experts, hidden).half()
        tensors[f"model.layers.{layer}.input_layernorm.weight"] = torch.ones(hidden, dtype=torch.float16)
    return tensors, full


@pytest.mark.parametrize("fmt", [(8, 8, 1), (16, 8, 1)])
def test_vocab_layers_and_experts_are_placed(fmt):
    b, d, m = fmt
    cfg = P.PQConfig(d, b, m, 1, lm_head_quantized=True)
    model = TinyMoEModel(cfg)
    tensors, full = moe_checkpoint(fmt)
    loaded, skipped = L.load_pq_linears(model, tensors.items())
    assert skipped == [n for n in tensors if "input_layernorm" in n]
    want = {n for n, _ in model.named_parameters()}
    assert loaded == want, sorted(want - loaded)
    # vocabulary: 100 rows of the checkpoint, 28 zero rows of padding (a zero scale -> a zero row)
    for mod, prefix in ((model.embed_tokens, "model.embed_tokens"), (model.lm_head, "lm_head")):
        c = full[prefix]
        assert mod.codes.shape[0] == 128
        assert torch.equal(mod.codes[:100], c["codes"]) and torch.equal(mod.scales[:100], c["scales"])
        assert torch.equal(mod.codebooks, c["codebooks"])
        assert not mod.codes[100:].any() and not mod.scales[100:].any()
    for li, layer in enumerate(model.layers):
        moe = layer.block_sparse_moe
        assert torch.equal(moe.gate_weight, tensors[f"model.layers.{li}.block_sparse_moe.gate.weight"])
        for e, expert in enumerate(moe.experts):
            for w in ("w1", "w2", "w3"):
                c = full[f"model.layers.{li}.block_sparse_moe.experts.{e}.{w}"]
                for key in ("codes", "codebooks", "scales"):
                    assert torch.equal(getattr(getattr(expert, w), key), c[key])


def test_tied_lm_head_and_mismatches():
    cfg = P.PQConfig(8, 8, 1, 1, lm_head_quantized=True)
    tied = TinyMoEModel(cfg, tie=True)
    assert "lm_head.codes" not in dict(tied.named_parameters())
    tensors, full = moe_checkpoint((8, 8, 1))
    loaded, skipped = L.load_pq_linears(tied, tensors.items())       # lm_head.* in the file: consumed, not placed
    assert not any("lm_head" in n for n in loaded) and not any("lm_head" in s for s in skipped)
    assert torch.equal(tied.lm_head.codes[:100], full["model.embed_tokens"]["codes"])
    # an expert index the model does not have, and a MoE tensor for a layer without a MoE block
    model = TinyMoEModel(cfg, experts=2)
    with pytest.raises(KeyError, match="experts.2"):
        L.load_pq_linears(model, tensors.items())
    model = TinyMoEModel(cfg)
    extra = {"model.layers.5.block_sparse_moe.experts.0.w1.codes": tensors["model.layers.0.block_sparse_moe.experts.0.w1.codes"]}
    with pytest.raises(KeyError, match="no MoE block"):
        L.load_pq_linears(model, extra.items())
    bad = {"model.layers.0.block_sparse_moe.gate.weight": torch.zeros(3, 128).half()}
    with pytest.raises(ValueError, match="shape"):
        L.load_pq_linears(model, bad.items())
    # a dense model (the Llama stack) still reports PQ vocabulary tensors as skipped
    stack = S.PQLlamaLinearStack(SHAPE, P.PQConfig(8, 8, 1, 1), torch.float16)
    _, skipped = L.load_pq_linears(stack, {"model.embed_tokens.codes": tensors["model.embed_tokens.codes"]}.items())
    assert skipped == ["model.embed_tokens.codes"]
