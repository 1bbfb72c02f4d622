"""PQ vocabulary layers without a GPU: the config switch, the parameters create_weights
registers for VocabParallelEmbedding / ParallelLMHead, the padded + sharded weight loader."""
import pytest
import torch

import vllm_pq_b200 as P
from oracle import pq_oracle as O


def test_config_switch_and_parameters():
    dense = P.PQConfig.from_config({"in_group_size": 8, "nbits_per_codebook": 8, "num_codebooks": 2, "out_group_size": 1})
    assert dense.lm_head_quantized is False
    with pytest.raises(NotImplementedError):          # as the reference: vocabulary layers stay dense (aqlm.py:213-217)
        P.VocabParallelEmbedding(1000, 64, quant_config=dense)
    cfg = P.PQConfig.from_config({"in_group_size": 8, "nbits_per_codebook": 8, "num_codebooks": 2, "out_group_size": 1,
                                  "lm_head_quantized": True})
    emb = P.VocabParallelEmbedding(1000, 64, params_dtype=torch.float16, quant_config=cfg)
    assert emb.num_embeddings_padded == 1024          # padded to a multiple of 64 (vocab_parallel_embedding.py:60-66)
    assert emb.codes.shape == (1024, 8, 2) and emb.codes.dtype == torch.int8
    assert emb.codebooks.shape == (2, 256, 1, 8) and emb.scales.shape == (1024, 1, 1, 1)
    assert callable(emb.quant_method.embedding)
    head = P.ParallelLMHead(1000, 64, params_dtype=torch.float16, quant_config=cfg).tie_weights(emb)
    assert head.codes is emb.codes
    with pytest.raises(RuntimeError):
        head(torch.zeros(1, dtype=torch.long))


def test_loader_pads_the_vocabulary_with_zero_scale_rows():
    cfg = P.PQConfig(8, 8, 1, 1, lm_head_quantized=True)
    emb = P.VocabParallelEmbedding(1000, 64, params_dtype=torch.float16, quant_config=cfg)
    c = O.make_case([1000], 64, 8, 8, 1, dtype=torch.float16, seed=5)
    for name in ("codes", "codebooks", "scales"):
        p = getattr(emb, name)
        p.weight_loader(p, c[name])
    assert torch.equal(emb.codes[:1000], c["codes"]) and torch.equal(emb.scales[:1000], c["scales"])
    assert float(emb.scales[1000:].abs().max()) == 0.0 and int(emb.codes[1000:].abs().max()) == 0
    assert torch.equal(emb.codebooks, c["codebooks"])


def test_ops_have_no_cpu_path():
    cfg = P.PQConfig(8, 8, 1, 1, lm_head_quantized=True)
    emb = P.VocabParallelEmbedding(128, 64, params_dtype=torch.float16, quant_config=cfg)
    with pytest.raises(RuntimeError):
        emb(torch.zeros(3, dtype=torch.long))


def test_mixtral_moe_block_parameters():
    """Per-expert PQ linears (mixtral_quant.py:58-139): three ReplicatedLinear per expert carrying the PQ triple,
    a dense router, all experts local at world size 1."""
    cfg = P.PQConfig(8, 8, 2, 1)
    moe = P.PQMixtralMoE(8, 2, 256, 512, params_dtype=torch.float16, quant_config=cfg)
    assert moe.expert_indicies == list(range(8)) and all(e is not None for e in moe.experts)
    e0 = moe.experts[0]
    assert tuple(e0.w1.codes.shape) == (512, 32, 2) and tuple(e0.w2.codes.shape) == (256, 64, 2)
    assert tuple(e0.w3.codebooks.shape) == (2, 256, 1, 8) and tuple(moe.gate_weight.shape) == (8, 256)
    names = {n for n, _ in moe.named_parameters()}
    assert "experts.3.w2.codes" in names and "experts.3.w2.scales" in names      # the checkpoint's tensor names
