"""Sharded create_weights and the weight loaders against the REFERENCE's own linear.py + aqlm.py
(SURVEY.md A13), every tensor-parallel rank simulated in one process.

oracle/ref_loader.load_reference_linear() executes the reference's layers/linear.py, parameter.py,
model_executor/utils.py, quantization/base_config.py and quantization/aqlm.py unmodified under stubs whose
``vllm.distributed`` answers with a settable (rank, world).  The same full-size checkpoint tensors go
through the reference layer's ``weight_loader`` and through ours; every parameter must come out with the
same shape, dtype, attributes and bits.  Only where /root/reference exists (the build container): the
gloo tests (tests/test_tp_gloo.py) cover the same logic against the oracle's arithmetic everywhere.
"""
import pytest
import torch

import vllm_pq_b200 as P
from oracle import pq_oracle as O
from oracle import ref_loader
from vllm_pq_b200 import distributed as D

pytestmark = pytest.mark.skipif(not ref_loader.available(), reason="/root/reference not present (GPU box)")

FORMATS = [(8, 8, 1), (8, 8, 2), (16, 8, 1), (12, 16, 1)]
HIDDEN, HEAD, HEADS, KV_HEADS, FFN = 256, 32, 8, 2, 512
ATTRS = ("input_dim", "output_dim", "packed_dim", "pack_factor", "is_metadata", "output_partition_sizes", "ignore_warning")


@pytest.fixture
def tp(monkeypatch):
    R = ref_loader.load_reference_linear()

    def set_tp(rank, world):
        R.set_tp(rank, world)
        monkeypatch.setattr(D, "_TP_RANK", rank)
        monkeypatch.setattr(D, "_TP_WORLD", world)
    yield R, set_tp
    R.set_tp(0, 1)


def same_layer(ours, theirs, what):
    mine, ref = dict(ours.named_parameters()), dict(theirs.named_parameters())
    assert set(mine) == set(ref), what
    for name, p in ref.items():
        q = mine[name]
        assert q.shape == p.shape and q.dtype == p.dtype, f"{what}.{name}: {tuple(q.shape)} {q.dtype} vs {tuple(p.shape)} {p.dtype}"
        assert torch.equal(q.data.view(torch.int16) if q.dtype.is_floating_point else q.data,
                           p.data.view(torch.int16) if p.dtype.is_floating_point else p.data), f"{what}.{name}"
        for a in ATTRS:
            assert getattr(q, a, None) == getattr(p, a, None), f"{what}.{name}.{a}"


def fill_garbage(layer, seed):
    g = torch.Generator().manual_seed(seed)
    for p in layer.parameters():
        if p.dtype.is_floating_point:
            p.data.copy_(torch.randn(p.shape, generator=g))
        else:
            p.data.copy_(torch.randint(-100, 100, p.shape, generator=g))


def build_pair(R, fmt, kind):
    b, d, m = fmt
    rcfg, ocfg = R.aqlm.AQLMConfig(d, b, m, 1), P.PQConfig(d, b, m, 1)
    kw = dict(bias=True, params_dtype=torch.float16)
    L = R.linear
    if kind == "qkv":
        return (P.QKVParallelLinear(HIDDEN, HEAD, HEADS, KV_HEADS, quant_config=ocfg, **kw),
                L.QKVParallelLinear(HIDDEN, HEAD, HEADS, KV_HEADS, quant_config=rcfg, **kw))
    if kind == "gate_up":
        return (P.MergedColumnParallelLinear(HIDDEN, [FFN, FFN], quant_config=ocfg, **kw),
                L.MergedColumnParallelLinear(HIDDEN, [FFN, FFN], quant_config=rcfg, **kw))
    if kind == "column":
        return (P.ColumnParallelLinear(HIDDEN, FFN, quant_config=ocfg, **kw),
                L.ColumnParallelLinear(HIDDEN, FFN, quant_config=rcfg, **kw))
    if kind == "row":
        return (P.RowParallelLinear(FFN, HIDDEN, quant_config=ocfg, **kw),
                L.RowParallelLinear(FFN, HIDDEN, quant_config=rcfg, **kw))
    return (P.ReplicatedLinear(HIDDEN, FFN, quant_config=ocfg, **kw),
            L.ReplicatedLinear(HIDDEN, FFN, quant_config=rcfg, **kw))


def load_both(ours, theirs, tensors, shard_id=None):
    for layer in (ours, theirs):
        params = dict(layer.named_parameters())
        for key, t in tensors.items():
            if shard_id is None:
                params[key].weight_loader(params[key], t.clone())
            else:
                params[key].weight_loader(params[key], t.clone(), shard_id)


def case(n, k, fmt, seed):
    b, d, m = fmt
    c = O.make_case([n], k, b, d, m, t=1, dtype=torch.float16, seed=seed, bias=True)
    return {key: c[key] for key in ("codes", "codebooks", "scales", "bias")}


@pytest.mark.parametrize("world", [1, 2, 4, 8])
@pytest.mark.parametrize("fmt", FORMATS, ids=["8x8x1", "8x8x2", "16x8x1", "12x16x1"])
def test_every_rank_loads_what_the_reference_loads(tp, fmt, world):
    R, set_tp = tp
    b, d, m = fmt
    for rank in range(world):
        set_tp(rank, world)
        # q / k / v and gate / up as separate checkpoint tensors (llama.py:371-378)
        ours, theirs = build_pair(R, fmt, "qkv")
        fill_garbage(ours, 1); fill_garbage(theirs, 2)
        for sid, n in (("q", HEADS * HEAD), ("k", KV_HEADS * HEAD), ("v", KV_HEADS * HEAD)):
            load_both(ours, theirs, case(n, HIDDEN, fmt, 20 + ord(sid)), sid)
        same_layer(ours, theirs, f"qkv rank {rank}/{world}")
        assert list(ours.output_partition_sizes) == list(theirs.output_partition_sizes)
        assert (ours.num_heads, ours.num_kv_heads, ours.num_kv_head_replicas) == \
               (theirs.num_heads, theirs.num_kv_heads, theirs.num_kv_head_replicas)

        ours, theirs = build_pair(R, fmt, "gate_up")
        fill_garbage(ours, 3); fill_garbage(theirs, 4)
        for sid in (0, 1):
            load_both(ours, theirs, case(FFN, HIDDEN, fmt, 30 + sid), sid)
        same_layer(ours, theirs, f"gate_up rank {rank}/{world}")

        for kind, n, k in (("column", FFN, HIDDEN), ("row", HIDDEN, FFN), ("replicated", FFN, HIDDEN)):
            if kind == "row" and (FFN // d) % world:
                continue
            ours, theirs = build_pair(R, fmt, kind)
            fill_garbage(ours, 5); fill_garbage(theirs, 6)
            load_both(ours, theirs, case(n, k, fmt, 40))
            same_layer(ours, theirs, f"{kind} rank {rank}/{world}")
            if kind == "row":
                assert ours.input_size_per_partition == theirs.input_size_per_partition
                # the bias is added by rank 0 only (linear.py:1206-1208)
                assert ours.skip_bias_add == theirs.skip_bias_add


@pytest.mark.parametrize("world", [1, 2, 4])
def test_fused_on_disk_checkpoints(tp, world):
    """qkv_proj / gate_up_proj already concatenated in the checkpoint: loaded_shard_id is None
    (linear.py:506-559, 890-944).  The codes and scales are split per shard by both; the codebooks
    (is_metadata, no output_dim) are copied whole."""
    R, set_tp = tp
    fmt = (8, 8, 1)
    for rank in range(world):
        set_tp(rank, world)
        ours, theirs = build_pair(R, fmt, "qkv")
        c = O.make_case([HEADS * HEAD, KV_HEADS * HEAD, KV_HEADS * HEAD], HIDDEN, 8, 8, 1, t=1, dtype=torch.float16, seed=50, bias=True)
        load_both(ours, theirs, {key: c[key] for key in ("codes", "codebooks", "scales", "bias")})
        same_layer(ours, theirs, f"fused qkv rank {rank}/{world}")
        ours, theirs = build_pair(R, fmt, "gate_up")
        c = O.make_case([FFN, FFN], HIDDEN, 8, 8, 1, t=1, dtype=torch.float16, seed=51, bias=True)
        load_both(ours, theirs, {key: c[key] for key in ("codes", "codebooks", "scales", "bias")})
        same_layer(ours, theirs, f"fused gate_up rank {rank}/{world}")


def test_construction_errors_match(tp):
    R, set_tp = tp
    set_tp(0, 3)
    cfg_r, cfg_o = R.aqlm.AQLMConfig(8, 8, 1, 1), P.PQConfig(8, 8, 1, 1)
    for make in (lambda L, c: L.ColumnParallelLinear(HIDDEN, 100, bias=False, params_dtype=torch.float16, quant_config=c),
                 lambda L, c: L.RowParallelLinear(FFN, HIDDEN, bias=False, params_dtype=torch.float16, quant_config=c)):
        with pytest.raises((AssertionError, ValueError)):
            make(R.linear, cfg_r)
        with pytest.raises((AssertionError, ValueError)):
            make(P, cfg_o)
    set_tp(0, 1)
    for cfg, mod in ((cfg_r, R.linear), (cfg_o, P)):
        with pytest.raises(ValueError):                  # aqlm.py:238-239: half precision only
            mod.ColumnParallelLinear(HIDDEN, FFN, bias=False, params_dtype=torch.float32, quant_config=cfg)


# ---- the quantised Mixtral block against the reference's own mixtral_quant.py ----------
@pytest.mark.parametrize("world", [1, 2, 4])
@pytest.mark.parametrize("fmt", [(8, 8, 1), (8, 8, 3)], ids=["8x8x1", "8x8x3"])
def test_mixtral_block_matches_reference_forward(tp, monkeypatch, fmt, world):
    """MixtralMoE.forward of the reference (mixtral_quant.py:133-160) runs on the CPU for formats outside
    its two kernel formats: its AQLM ``apply`` falls to generic_dequantize_gemm (aqlm.py:366-374).  Our
    block with ``pq::gemm`` replaced by the oracle's restatement of that function must give the same
    values (bit for bit, up to the sign of zeros): routing, expert split between ranks, weighting,
    accumulation order, skipped experts."""
    from transformers import MixtralConfig

    from vllm_pq_b200 import _custom_ops as ops
    R, set_tp = tp
    b, d, m = fmt
    hidden, ffn, n_exp, top_k = 128, 256, 4, 2
    monkeypatch.setattr(ops, "pq_gemm", lambda x, codes, codebooks, scales, parts, bias, algo=0:
                        O.generic_dequantize_gemm(x, codes, codebooks, scales, parts, bias))
    R.set_all_reduce(lambda t: t)                        # each simulated rank's partial sum
    monkeypatch.setattr(D, "tensor_model_parallel_all_reduce", lambda t: t)
    import vllm_pq_b200.moe as moe_mod
    monkeypatch.setattr(moe_mod, "tensor_model_parallel_all_reduce", lambda t: t)
    gen = torch.Generator().manual_seed(3)
    gate = torch.randn(n_exp, hidden, generator=gen).half()
    weights = {}
    for e in range(n_exp):
        for name, n, k in (("w1", ffn, hidden), ("w3", ffn, hidden), ("w2", hidden, ffn)):
            c = O.make_case([n], k, b, d, m, dtype=torch.float16, seed=200 + 10 * e + ord(name[1]))
            c["codebooks"] = (c["codebooks"].float() / (k * m) ** 0.5).half()
            weights[(e, name)] = c
    try:
        for rank in range(world):
            set_tp(rank, world)
            torch.set_default_dtype(torch.float16)
            try:
                theirs = R.mixtral.MixtralMoE(MixtralConfig(hidden_size=hidden, intermediate_size=ffn, num_local_experts=n_exp,
                                                            num_experts_per_tok=top_k), R.aqlm.AQLMConfig(d, b, m, 1))
            finally:
                torch.set_default_dtype(torch.float32)
            ours = P.PQMixtralMoE(n_exp, top_k, hidden, ffn, params_dtype=torch.float16, quant_config=P.PQConfig(d, b, m, 1))
            assert ours.expert_indicies == theirs.expert_indicies
            assert [e is None for e in ours.experts] == [e is None for e in theirs.experts]
            theirs.gate.weight.data.copy_(gate)
            ours.gate_weight.data.copy_(gate)
            for e in ours.expert_indicies:
                for name in ("w1", "w2", "w3"):
                    for layer in (getattr(ours.experts[e], name), getattr(theirs.experts[e], name)):
                        for key in ("codes", "codebooks", "scales"):
                            p = getattr(layer, key)
                            p.weight_loader(p, weights[(e, name)][key])
            for t in (1, 5, 40):
                x = torch.randn(t, hidden, generator=gen).half()
                with torch.no_grad():
                    y_ref = theirs(x.clone())
                    y = ours(x.clone())
                assert y.shape == y_ref.shape and y.dtype == y_ref.dtype
                # equal as numbers: identical bits except the sign of a zero - an expert no token chose is
                # skipped here (+0) and multiplied by a zero weight there (-0 where its output is negative)
                assert torch.equal(y, y_ref), f"rank {rank}/{world}, {t} tokens"
                nz = y_ref != 0
                assert torch.equal(y.view(torch.int16)[nz], y_ref.view(torch.int16)[nz])
    finally:
        R.set_all_reduce(None)


# ---- layer forward against the reference's own forward (generic formats run on the CPU there) ---------
@pytest.mark.parametrize("world", [1, 2, 4])
@pytest.mark.parametrize("fmt", [(8, 8, 1), (4, 8, 2)], ids=["8x8x1", "4x8x2"])
def test_layer_forward_matches_reference_forward(tp, monkeypatch, fmt, world):
    """ColumnParallel / Merged / QKV / RowParallel / Replicated ``forward`` (linear.py:285-295,410-435,
    1193-1221): what goes to ``apply`` (the input chunk of a non-parallel row input, bias on rank 0 only,
    skip_bias_add) and what comes back (output, output_bias).  ``pq::gemm`` is replaced by the oracle's
    generic_dequantize_gemm - bit-equal to the reference's (tests/test_oracle_golden.py) - and the
    all-reduce by the identity on both sides, so every simulated rank's partial result is compared."""
    from vllm_pq_b200 import _custom_ops as ops
    R, set_tp = tp
    b, d, m = fmt
    monkeypatch.setattr(ops, "pq_gemm", lambda x, codes, codebooks, scales, parts, bias, algo=0:
                        O.generic_dequantize_gemm(x, codes, codebooks, scales, parts, bias))
    monkeypatch.setattr(D, "tensor_model_parallel_all_reduce", lambda t: t)
    import vllm_pq_b200.linear as lin_mod
    monkeypatch.setattr(lin_mod, "tensor_model_parallel_all_reduce", lambda t: t)
    R.set_all_reduce(lambda t: t)
    gen = torch.Generator().manual_seed(5)
    rcfg, ocfg = R.aqlm.AQLMConfig(d, b, m, 1), P.PQConfig(d, b, m, 1)
    kw = dict(params_dtype=torch.float16)

    def check(ours, theirs, x, what):
        with torch.no_grad():
            y_ref, b_ref = theirs(x.clone())
            y, b_out = ours(x.clone())
        assert y.shape == y_ref.shape and torch.equal(y.view(torch.int16), y_ref.view(torch.int16)), what
        assert (b_out is None) == (b_ref is None), what
        if b_ref is not None:
            assert torch.equal(b_out, b_ref), what

    try:
        for rank in range(world):
            set_tp(rank, world)
            x = torch.randn(2, 3, HIDDEN, generator=gen).half()
            xr_full = torch.randn(2, 3, FFN, generator=gen).half()
            for bias, skip in ((True, False), (True, True), (False, False)):
                tag = f"rank {rank}/{world} bias={bias} skip_bias_add={skip}"
                pairs = [
                    ("qkv", P.QKVParallelLinear(HIDDEN, HEAD, HEADS, KV_HEADS, bias=bias, skip_bias_add=skip, quant_config=ocfg, **kw),
                     R.linear.QKVParallelLinear(HIDDEN, HEAD, HEADS, KV_HEADS, bias=bias, skip_bias_add=skip, quant_config=rcfg, **kw), x),
                    ("gate_up", P.MergedColumnParallelLinear(HIDDEN, [FFN, FFN], bias=bias, skip_bias_add=skip, quant_config=ocfg, **kw),
                     R.linear.MergedColumnParallelLinear(HIDDEN, [FFN, FFN], bias=bias, skip_bias_add=skip, quant_config=rcfg, **kw), x),
                    ("replicated", P.ReplicatedLinear(HIDDEN, FFN, bias=bias, skip_bias_add=skip, quant_config=ocfg, **kw),
                     R.linear.ReplicatedLinear(HIDDEN, FFN, bias=bias, skip_bias_add=skip, quant_config=rcfg, **kw), x),
                ]
                for parallel_in in (True, False):
                    xr = xr_full[..., rank * (FFN // world):(rank + 1) * (FFN // world)].contiguous() if parallel_in else xr_full
                    pairs.append((f"row(input_is_parallel={parallel_in})",
                                  P.RowParallelLinear(FFN, HIDDEN, bias=bias, skip_bias_add=skip, input_is_parallel=parallel_in,
                                                      quant_config=ocfg, **kw),
                                  R.linear.RowParallelLinear(FFN, HIDDEN, bias=bias, skip_bias_add=skip,
                                                             input_is_parallel=parallel_in, quant_config=rcfg, **kw), xr))
                for name, ours, theirs, xin in pairs:
                    seeds = {"qkv": 61, "gate_up": 62, "replicated": 63}.get(name, 64)
                    fused = {"qkv": [("q", HEADS * HEAD), ("k", KV_HEADS * HEAD), ("v", KV_HEADS * HEAD)],
                             "gate_up": [(0, FFN), (1, FFN)]}.get(name)
                    if fused:
                        for i, (sid, n) in enumerate(fused):
                            c = case(n, HIDDEN, fmt, seeds * 10 + i)
                            c["codebooks"] = (c["codebooks"].float() / (HIDDEN * m) ** 0.5).half()
                            load_both(ours, theirs, {k_: v for k_, v in c.items() if k_ != "bias" or bias}, sid)
                    else:
                        n, k = (HIDDEN, FFN) if name.startswith("row") else (FFN, HIDDEN)
                        c = case(n, k, fmt, seeds)
                        c["codebooks"] = (c["codebooks"].float() / (k * m) ** 0.5).half()
                        load_both(ours, theirs, {k_: v for k_, v in c.items() if k_ != "bias" or bias})
                    check(ours, theirs, xin, f"{name} {tag}")
    finally:
        R.set_all_reduce(None)


# ---- vocabulary layers against the reference's own vocab_parallel_embedding.py ---------
@pytest.mark.parametrize("world", [1, 2, 4])
@pytest.mark.parametrize("vocab", [100, 320, 32000])
def test_vocab_layers_match_reference(tp, monkeypatch, vocab, world):
    """The reference keeps its vocabulary layers dense under AQLM (aqlm.py:213-217), but the classes take
    any method that implements ``embedding`` (vocab_parallel_embedding.py:229-246).  Here they get the
    reference's AQLM method plus an ``embedding`` that looks rows up in the reference's own
    dequantize_weight, so padding, shard ranges, ``create_weights``, ``weight_loader`` (zero padding rows)
    and the masked lookup of ``forward`` are the reference's code; ours runs with ``pq::embedding`` /
    ``pq::gemm`` replaced by the oracle."""
    import torch.nn.functional as F

    from vllm_pq_b200 import _custom_ops as ops
    from vllm_pq_b200 import embedding as emb_mod
    R, set_tp = tp
    b, d, m, hidden = 8, 8, 1, 64

    class Method(R.aqlm.AQLMLinearMethod):
        def embedding(self, layer, input_):
            w = R.aqlm.dequantize_weight(R.aqlm.unpack_int_data(layer.codes, b), layer.codebooks, layer.scales)
            return F.embedding(input_, w)

    class Config(R.aqlm.AQLMConfig):
        def get_quant_method(self, layer, prefix):
            return Method(self)

    def cpu_embedding(ids, codes, codebooks, scales):
        w = O.dequantize_partitions(codes, codebooks, [codes.shape[0]], scales)
        out = torch.zeros(ids.shape + (w.shape[1],), dtype=w.dtype)
        ok = (ids >= 0) & (ids < w.shape[0])
        out[ok] = w[ids[ok]]
        return out

    monkeypatch.setattr(ops, "pq_embedding", cpu_embedding)
    monkeypatch.setattr(ops, "pq_gemm", lambda x, codes, codebooks, scales, parts, bias, algo=0:
                        O.generic_dequantize_gemm(x, codes, codebooks, scales, parts, bias))
    monkeypatch.setattr(emb_mod, "tensor_model_parallel_all_reduce", lambda t: t)
    R.set_all_reduce(lambda t: t)
    c = O.make_case([vocab], hidden, b, d, m, t=3, dtype=torch.float16, seed=77, bias=True)
    gen = torch.Generator().manual_seed(8)
    ids = torch.randint(0, vocab, (2, 9), generator=gen)
    ids[0, :3] = torch.tensor([0, vocab - 1, vocab // 2])
    try:
        for rank in range(world):
            set_tp(rank, world)
            kw = dict(params_dtype=torch.float16)
            for cls_o, cls_r in ((P.VocabParallelEmbedding, R.vocab.VocabParallelEmbedding),
                                 (P.ParallelLMHead, R.vocab.ParallelLMHead)):
                extra = dict(bias=True) if cls_o is P.ParallelLMHead else {}
                ours = cls_o(vocab, hidden, quant_config=P.PQConfig(d, b, m, 1, lm_head_quantized=True), **extra, **kw)
                theirs = cls_r(vocab, hidden, quant_config=Config(d, b, m, 1), **extra, **kw)
                fill_garbage(ours, 1); fill_garbage(theirs, 2)
                # vocab_parallel_embedding.py:376-380 reads `param.packed_factor` inside an assert where plain
                # Parameters carry `pack_factor` (the AQLM scales: packed_dim == output_dim, factor 1) - a path
                # the reference never took, its AQLM vocabulary layers being dense.  Give it the name it reads.
                theirs.scales.packed_factor = theirs.scales.pack_factor
                tensors = {k: c[k] for k in ("codes", "codebooks", "scales")}
                if extra:
                    tensors["bias"] = c["bias"]
                load_both(ours, theirs, tensors)
                same_layer(ours, theirs, f"{cls_o.__name__} rank {rank}/{world}")
                assert ours.num_embeddings_padded == theirs.num_embeddings_padded
                assert ours.num_embeddings_per_partition == theirs.num_embeddings_per_partition
                si = theirs.shard_indices                    # the reference clamps its range to the real vocabulary
                assert (min(ours.vocab_start, vocab), min(ours.vocab_end, vocab)) == \
                       (si.org_vocab_start_index, si.org_vocab_end_index)
                if cls_o is P.VocabParallelEmbedding:
                    with torch.no_grad():
                        y_ref, y = theirs(ids.clone()), ours(ids.clone())
                    assert torch.equal(y, y_ref), f"lookup rank {rank}/{world}"      # +-0 in masked rows
                    nz = y_ref != 0
                    assert torch.equal(y.view(torch.int16)[nz], y_ref.view(torch.int16)[nz])
                else:
                    with pytest.raises(RuntimeError):
                        ours(ids)
                    with pytest.raises(RuntimeError):
                        theirs(ids)
                    # logits_processor.py:_get_logits: lm_head.quant_method.apply(lm_head, h, bias)
                    with torch.no_grad():
                        lg_ref = theirs.quant_method.apply(theirs, c["x"], bias=theirs.bias)
                        lg = ours.compute_logits(c["x"])
                    assert torch.equal(lg.view(torch.int16), lg_ref.view(torch.int16))
    finally:
        R.set_all_reduce(None)


# ---- the drop-in: OUR method behind the REFERENCE's plugin API, inside the reference's own layers -------
def test_pq_method_plugs_into_the_reference_tree(tp, monkeypatch):
    """What INTEGRATION.md tells a maintainer to do, done against the reference's real modules:
    ``make_classes`` mounts the PQ config / method bodies on the reference's QuantizationConfig and
    LinearMethodBase, ``register_quantization_config`` (quantization/__init__.py:39-73) registers the
    result, and the reference's own QKV / merged / row-parallel layers build, load (their loaders) and
    run (their forward) with it - giving the parameters and, with ``pq::gemm`` replaced by the oracle,
    the outputs that the reference's AQLM method gives in the same layers."""
    from vllm_pq_b200 import _custom_ops as ops
    from vllm_pq_b200.pq import make_classes
    R, set_tp = tp
    cfg_cls, method_cls = make_classes(R.base_config.QuantizationConfig, R.linear.LinearMethodBase,
                                       lambda layer: isinstance(layer, R.linear.LinearBase), name="pq")
    assert issubclass(cfg_cls, R.base_config.QuantizationConfig) and issubclass(method_cls, R.linear.LinearMethodBase)
    if "pq" not in R.registry.QUANTIZATION_METHODS:
        assert R.registry.register_quantization_config("pq")(cfg_cls) is cfg_cls
    assert "pq" in R.registry.QUANTIZATION_METHODS and "pq" in R.registry._CUSTOMIZED_METHOD_TO_QUANT_CONFIG
    with pytest.raises(ValueError, match="already exists"):
        R.registry.register_quantization_config("pq")(cfg_cls)
    with pytest.raises(ValueError, match="already exists"):
        R.registry.register_quantization_config("aqlm")(cfg_cls)
    hf = {"in_group_size": 8, "nbits_per_codebook": 8, "num_codebooks": 1, "out_group_size": 1}
    pq_cfg, aqlm_cfg = cfg_cls.from_config(dict(hf)), R.aqlm.AQLMConfig.from_config(dict(hf))
    assert pq_cfg.get_name() == "pq" and cfg_cls.get_config_filenames() == R.aqlm.AQLMConfig.get_config_filenames()
    assert pq_cfg.get_quant_method(torch.nn.Linear(8, 8), prefix="") is None          # aqlm.py:213-217

    monkeypatch.setattr(ops, "pq_gemm", lambda x, codes, codebooks, scales, parts, bias, algo=0:
                        O.generic_dequantize_gemm(x, codes, codebooks, scales, parts, bias))
    R.set_all_reduce(lambda t: t)
    gen = torch.Generator().manual_seed(12)
    L = R.linear
    try:
        for world in (1, 2, 4):
            for rank in range(world):
                set_tp(rank, world)
                kw = dict(bias=True, params_dtype=torch.float16)
                layers = [
                    (lambda c: L.QKVParallelLinear(HIDDEN, HEAD, HEADS, KV_HEADS, quant_config=c, **kw),
                     [("q", HEADS * HEAD), ("k", KV_HEADS * HEAD), ("v", KV_HEADS * HEAD)], HIDDEN, HIDDEN),
                    (lambda c: L.MergedColumnParallelLinear(HIDDEN, [FFN, FFN], quant_config=c, **kw),
                     [(0, FFN), (1, FFN)], HIDDEN, HIDDEN),
                    (lambda c: L.RowParallelLinear(FFN, HIDDEN, quant_config=c, **kw), [(None, HIDDEN)], FFN, FFN // world),
                ]
                for make, shards, k, k_in in layers:
                    with_pq, with_aqlm = make(pq_cfg), make(aqlm_cfg)
                    assert type(with_pq.quant_method) is method_cls
                    for i, (sid, n) in enumerate(shards):
                        c = case(n, k, (8, 8, 1), 300 + i)
                        c["codebooks"] = (c["codebooks"].float() / k ** 0.5).half()
                        load_both(with_pq, with_aqlm, c, sid)
                    same_layer(with_pq, with_aqlm, f"{type(with_pq).__name__} rank {rank}/{world}")
                    x = torch.randn(7, k_in, generator=gen).half()
                    with torch.no_grad():
                        y, _ = with_pq(x.clone())
                        y_ref, _ = with_aqlm(x.clone())
                    assert torch.equal(y.view(torch.int16), y_ref.view(torch.int16))
    finally:
        R.set_all_reduce(None)


# ---- the checkpoint walk against the reference's own LlamaModel.load_weights -----------
@pytest.mark.parametrize("world", [1, 2, 4])
@pytest.mark.parametrize("fmt", [(8, 8, 1), (8, 8, 2), (16, 8, 1)], ids=["8x8x1", "8x8x2", "16x8x1"])
def test_loader_places_what_the_reference_load_weights_places(tp, tmp_path, fmt, world):
    """loader.load_pq_linears on our stack against LlamaModel.load_weights (llama.py:369-433, run unmodified as
    a plain function) on a module tree of the reference's own layers with its AQLM method: the same
    checkpoint tensors, every simulated rank, the same set of loaded parameters and the same bits."""
    from tests.test_loader_cpu import SHAPE, make_checkpoint
    from vllm_pq_b200 import loader as L
    from vllm_pq_b200 import stack as S
    R, set_tp = tp
    b, d, m = fmt
    make_checkpoint(str(tmp_path), fmt)
    weights = list(L.safetensors_weights_iterator(str(tmp_path)))
    hd = SHAPE.head_dim

    class Block(torch.nn.Module):
        pass

    for rank in range(world):
        set_tp(rank, world)
        ours = S.PQLlamaLinearStack(SHAPE, P.PQConfig(d, b, m, 1), torch.float16)
        loaded, skipped = L.load_pq_linears(ours, weights)
        rcfg = R.aqlm.AQLMConfig(d, b, m, 1)
        kw = dict(bias=False, params_dtype=torch.float16, quant_config=rcfg)
        tree = Block()
        tree.quant_config = rcfg
        tree.layers = torch.nn.ModuleList()
        for _ in range(SHAPE.layers):
            layer, layer.self_attn, layer.mlp = Block(), Block(), Block()
            layer.self_attn.qkv_proj = R.linear.QKVParallelLinear(SHAPE.hidden, hd, SHAPE.heads, SHAPE.kv_heads, **kw)
            layer.self_attn.o_proj = R.linear.RowParallelLinear(SHAPE.heads * hd, SHAPE.hidden, **kw)
            layer.mlp.gate_up_proj = R.linear.MergedColumnParallelLinear(SHAPE.hidden, [SHAPE.intermediate] * 2, **kw)
            layer.mlp.down_proj = R.linear.RowParallelLinear(SHAPE.intermediate, SHAPE.hidden, **kw)
            tree.layers.append(layer)
        fill_garbage(tree, 7)
        # what LlamaForCausalLM hands LlamaModel.load_weights: the "model." prefix stripped (AutoWeightsLoader);
        # only the tensors of modules this tree has (no norms / embeddings here - ours reports those as skipped)
        pq_names = {n for n, _ in weights if n not in skipped}
        ref_loaded = R.llama.LlamaModel.load_weights(
            tree, ((n[len("model."):], t) for n, t in weights if n in pq_names or "rotary_emb" in n))
        strip = lambda n: n.replace(".self_attn.", ".").replace(".mlp.", ".")          # noqa: E731
        assert {strip(n) for n in ref_loaded} == loaded
        assert all("rotary_emb" in s or "layernorm" in s or "embed_tokens" in s or "lm_head" in s for s in skipped)
        mine, ref = dict(ours.named_parameters()), dict(tree.named_parameters())
        assert {strip(n) for n in ref} == set(mine)
        for name, p in ref.items():
            q = mine[strip(name)]
            assert q.shape == p.shape and q.dtype == p.dtype, name
            assert torch.equal(q.data.view(torch.int16) if q.dtype.is_floating_point else q.data,
                               p.data.view(torch.int16) if p.dtype.is_floating_point else p.data), f"{name} rank {rank}/{world}"


# ---- the MLP with the activation on down_proj's input against the reference's LlamaMLP ----
@pytest.mark.parametrize("world", [1, 2, 4])
@pytest.mark.parametrize("bias", [False, True])
def test_mlp_with_input_activation_matches_reference_llama_mlp(tp, monkeypatch, world, bias):
    """LlamaMLP.forward (llama.py:89-93): gate_up_proj -> SiluAndMul -> down_proj, the reference's own class on
    the CPU (generic format).  Ours: ``down_proj(gate_up, input_act="silu_and_mul")``, the activation taken by
    ``pq::gemm_silu_mul`` - here the oracle's SiLU*mul (bit-equal to the reference's, test_oracle_golden.py)
    followed by its generic_dequantize_gemm.  Every simulated rank's partial result, bit for bit."""
    from vllm_pq_b200 import _custom_ops as ops
    R, set_tp = tp
    fmt = (8, 8, 1)
    b, d, m = fmt

    def cpu_gemm(x, codes, codebooks, scales, parts, bias, algo=0):
        return O.generic_dequantize_gemm(x, codes, codebooks, scales, parts, bias)

    monkeypatch.setattr(ops, "pq_gemm", cpu_gemm)
    monkeypatch.setattr(ops, "pq_gemm_silu_mul", lambda x, *a, **k: cpu_gemm(O.silu_and_mul(x), *a, **k))
    import vllm_pq_b200.linear as lin_mod
    monkeypatch.setattr(lin_mod, "tensor_model_parallel_all_reduce", lambda t: t)
    monkeypatch.setattr(D, "tensor_model_parallel_all_reduce", lambda t: t)
    R.set_all_reduce(lambda t: t)
    gen = torch.Generator().manual_seed(21)
    gate, up, down = case(FFN, HIDDEN, fmt, 401), case(FFN, HIDDEN, fmt, 402), case(HIDDEN, FFN, fmt, 403)
    for c, k in ((gate, HIDDEN), (up, HIDDEN), (down, FFN)):
        c["codebooks"] = (c["codebooks"].float() / k ** 0.5).half()
    try:
        for rank in range(world):
            set_tp(rank, world)
            torch.set_default_dtype(torch.float16)
            try:
                theirs = R.llama.LlamaMLP(HIDDEN, FFN, "silu", quant_config=R.aqlm.AQLMConfig(d, b, m, 1), bias=bias)
            finally:
                torch.set_default_dtype(torch.float32)
            kw = dict(bias=bias, params_dtype=torch.float16, quant_config=P.PQConfig(d, b, m, 1))
            gu, dn = P.MergedColumnParallelLinear(HIDDEN, [FFN, FFN], **kw), P.RowParallelLinear(FFN, HIDDEN, **kw)
            pick = (lambda c: c) if bias else (lambda c: {k_: v for k_, v in c.items() if k_ != "bias"})
            load_both(gu, theirs.gate_up_proj, pick(gate), 0)
            load_both(gu, theirs.gate_up_proj, pick(up), 1)
            load_both(dn, theirs.down_proj, pick(down))
            x = torch.randn(5, HIDDEN, generator=gen).half()
            with torch.no_grad():
                y_ref = theirs(x.clone())
                h, _ = gu(x.clone())
                y, _ = dn(h, input_act="silu_and_mul")
                y_sep, _ = dn(O.silu_and_mul(h))                  # the activation as its own step: same bits
            assert torch.equal(y.view(torch.int16), y_ref.view(torch.int16)), f"rank {rank}/{world}"
            assert torch.equal(y_sep.view(torch.int16), y_ref.view(torch.int16))
    finally:
        R.set_all_reduce(None)
