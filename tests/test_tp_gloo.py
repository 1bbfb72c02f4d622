"""Tensor-parallel host logic on the CPU: two processes, gloo backend.

What runs here is everything of the TP path that is not a kernel: sharded ``create_weights``
(linear.py:324-333,759-792,1102-1129 in the reference), the v1 weight loaders narrowing full
checkpoint tensors to this rank's shard (linear.py:370-408,494-640,873-1058,1144-1180), and the
reduction hook.  The arithmetic of each rank's shard is done by the ORACLE (tests may use it):
sum over ranks of the row-parallel partials and concatenation of the column-parallel shards
must reproduce the unsharded oracle result.
"""
import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port() -> int:
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank: int, world: int, port: int, fmt, ret):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import vllm_pq_b200 as P
        from oracle import pq_oracle as O
        P.init_tensor_parallel()
        assert P.get_tensor_model_parallel_world_size() == world
        assert P.get_tensor_model_parallel_rank() == rank
        b, d, m = fmt
        cfg = P.PQConfig(d, b, m, 1)
        hidden, head, heads, kv_heads, inter = 256, 32, 8, 2, 512
        dt = torch.float16

        # ---- full (unsharded) checkpoint tensors, identical on both ranks ----
        def full(parts, k, seed):
            return O.make_case(parts, k, b, d, m, t=3, dtype=dt, seed=seed)
        q, kk, v = full([heads * head], hidden, 1), full([kv_heads * head], hidden, 2), full([kv_heads * head], hidden, 3)
        o = full([hidden], heads * head, 4)
        gate, up = full([inter], hidden, 5), full([inter], hidden, 6)
        down = full([hidden], inter, 7)

        kw = dict(bias=False, params_dtype=dt, quant_config=cfg)
        qkv_l = P.QKVParallelLinear(hidden, head, heads, kv_heads, **kw)
        o_l = P.RowParallelLinear(heads * head, hidden, **kw)
        gu_l = P.MergedColumnParallelLinear(hidden, [inter, inter], **kw)
        down_l = P.RowParallelLinear(inter, hidden, **kw)

        # shapes per rank
        g = hidden // d
        assert tuple(qkv_l.codes.shape) == ((heads + 2 * kv_heads) * head // world, g, m)
        assert tuple(qkv_l.codebooks.shape) == (3 * m, 1 << b, 1, d)
        assert tuple(o_l.codes.shape) == (hidden, heads * head // d // world, m)
        assert tuple(o_l.scales.shape) == (hidden, 1, 1, 1)
        assert tuple(gu_l.codes.shape) == (2 * inter // world, g, m)
        assert tuple(down_l.codes.shape) == (hidden, inter // d // world, m)

        # ---- the unmodified-loader path: full tensors in, this rank's shard out ----
        for name in ("codes", "codebooks", "scales"):
            for sid, c in (("q", q), ("k", kk), ("v", v)):
                p = getattr(qkv_l, name)
                p.weight_loader(p, c[name], sid)
            for sid, c in ((0, gate), (1, up)):
                p = getattr(gu_l, name)
                p.weight_loader(p, c[name], sid)
            for layer, c in ((o_l, o), (down_l, down)):
                p = getattr(layer, name)
                p.weight_loader(p, c[name])

        def rows(c, lo, hi):
            return c["codes"][lo:hi], c["scales"][lo:hi]
        qs, ks = heads * head // world, kv_heads * head // world
        exp_codes = torch.cat([q["codes"][rank * qs:(rank + 1) * qs], kk["codes"][rank * ks:(rank + 1) * ks],
                               v["codes"][rank * ks:(rank + 1) * ks]])
        assert torch.equal(qkv_l.codes.data, exp_codes)
        assert torch.equal(qkv_l.codebooks.data, torch.cat([q["codebooks"], kk["codebooks"], v["codebooks"]]))
        gs = (heads * head // d) // world
        assert torch.equal(o_l.codes.data, o["codes"][:, rank * gs:(rank + 1) * gs])
        assert torch.equal(o_l.scales.data, o["scales"])                      # replicated for row-parallel
        isz = inter // world
        assert torch.equal(gu_l.codes.data, torch.cat([gate["codes"][rank * isz:(rank + 1) * isz],
                                                       up["codes"][rank * isz:(rank + 1) * isz]]))

        # ---- TP math with the oracle as the per-shard compute ----
        def shard_linear(layer, x):
            parts = list(getattr(layer.codebooks, "output_partition_sizes"))
            return O.linear_fp32(x, layer.codes.data, layer.codebooks.data, layer.scales.data, parts, None)

        x = q["x"]                                                           # [3, hidden]
        # column-parallel: concatenating the ranks' outputs (per partition) gives the full output
        y_loc = shard_linear(qkv_l, x).to(torch.float32)
        gathered = [torch.empty_like(y_loc) for _ in range(world)]
        dist.all_gather(gathered, y_loc)
        full_q = O.linear_fp32(x, q["codes"], q["codebooks"], q["scales"], [heads * head], None).float()
        got_q = torch.cat([gt[:, :qs] for gt in gathered], dim=1)
        torch.testing.assert_close(got_q, full_q, rtol=1e-5, atol=1e-5)
        full_k = O.linear_fp32(x, kk["codes"], kk["codebooks"], kk["scales"], [kv_heads * head], None).float()
        got_k = torch.cat([gt[:, qs:qs + ks] for gt in gathered], dim=1)
        torch.testing.assert_close(got_k, full_k, rtol=1e-5, atol=1e-5)

        # row-parallel: each rank multiplies its K-slice; the hook sums the partials
        xin = o["x"]                                                         # [3, heads*head]
        chunk = xin.shape[1] // world
        part = shard_linear(o_l, xin[:, rank * chunk:(rank + 1) * chunk]).float()
        total = P.tensor_model_parallel_all_reduce(part.clone())
        full_o = O.linear_fp32(xin, o["codes"], o["codebooks"], o["scales"], [hidden], None).float()
        torch.testing.assert_close(total, full_o, rtol=1e-4, atol=1e-4)
        # the hook is out-of-place-or-in-place on a fresh tensor, identical on every rank
        both = [torch.empty_like(total) for _ in range(world)]
        dist.all_gather(both, total)
        assert torch.equal(both[0], both[1])

        # create_weights refuses a K shard that is not a whole number of groups (aqlm.py:240-244)
        with pytest.raises(ValueError):
            P.RowParallelLinear(d * 3, hidden, **kw)                          # 3 groups over 2 ranks
        ret[rank] = "ok"
    finally:
        P.destroy_tensor_parallel()
        dist.destroy_process_group()


@pytest.mark.parametrize("fmt", [(8, 8, 1), (8, 8, 2), (16, 8, 1)])
def test_tp2_sharding_loaders_and_reduce(fmt):
    world = 2
    ctx = mp.get_context("spawn")
    with ctx.Manager() as mgr:
        ret = mgr.dict()
        port = _free_port()
        procs = [ctx.Process(target=_worker, args=(r, world, port, fmt, ret)) for r in range(world)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(timeout=240)
        for p in procs:
            if p.is_alive():
                p.kill()
                pytest.fail("TP worker timed out")
            assert p.exitcode == 0, f"TP worker exit code {p.exitcode}"
        assert dict(ret) == {0: "ok", 1: "ok"}


# ---- the fused-chain hooks and the vocabulary layers (N4) at world size 2 -------------
# On the CPU the kernels do not exist, so the per-shard arithmetic is the ORACLE's, patched in for the
# ops; what is under test is the host logic of the separate-launch form: all-reduce BEFORE the
# residual add, bias on rank 0 only, sum of squares of the reduced stream, RMSNorm from that sum on
# the next column-parallel layer, masked vocabulary-parallel lookup + all-reduce.
def _worker_fused(rank: int, world: int, port: int, ret):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import vllm_pq_b200 as P
        from oracle import pq_oracle as O
        from vllm_pq_b200 import _custom_ops as ops
        P.init_tensor_parallel()

        def cpu_gemm(x, codes, codebooks, scales, parts, bias, algo=0):
            return O.linear_fp32(x, codes, codebooks, scales, parts, bias).to(x.dtype)

        def cpu_gemm_silu(x, codes, codebooks, scales, parts, bias, algo=0):
            return cpu_gemm(O.silu_and_mul(x), codes, codebooks, scales, parts, bias)

        def cpu_norm(x, residual, weight, eps):
            out, z = O.fused_add_rms_norm(x, residual, weight, eps)
            x.copy_(out)
            if residual is not None:
                residual.copy_(z)

        def cpu_embedding(ids, codes, codebooks, scales):
            w = O.dequantize_partitions(codes, codebooks, [codes.shape[0]], scales)
            out = torch.zeros(ids.shape + (w.shape[1],), dtype=w.dtype)
            ok = (ids >= 0) & (ids < w.shape[0])
            out[ok] = w[ids[ok]]
            return out
        ops.pq_gemm, ops.pq_gemm_silu_mul, ops.fused_add_rms_norm, ops.pq_embedding = cpu_gemm, cpu_gemm_silu, cpu_norm, cpu_embedding

        dt, eps = torch.float16, 1e-5
        cfg = P.PQConfig(8, 8, 1, 1, lm_head_quantized=True)
        hidden, inter, vocab = 256, 512, 300
        kw = dict(bias=True, params_dtype=dt, quant_config=cfg)
        o_full = O.make_case([hidden], hidden, 8, 8, 1, t=3, dtype=dt, seed=11, bias=True)
        gate, up = (O.make_case([inter], hidden, 8, 8, 1, t=3, dtype=dt, seed=s) for s in (12, 13))
        down_full = O.make_case([hidden], inter, 8, 8, 1, t=3, dtype=dt, seed=14)
        for c in (o_full, gate, up, down_full):
            c["codebooks"] = (c["codebooks"].float() / (c["codes"].shape[1] * 8) ** 0.5).to(dt)
        o_l = P.RowParallelLinear(hidden, hidden, **kw)
        gu_l = P.MergedColumnParallelLinear(hidden, [inter, inter], bias=False, params_dtype=dt, quant_config=cfg)
        down_l = P.RowParallelLinear(inter, hidden, bias=False, params_dtype=dt, quant_config=cfg)
        for name in ("codes", "codebooks", "scales"):
            p = getattr(o_l, name); p.weight_loader(p, o_full[name])
            p = getattr(down_l, name); p.weight_loader(p, down_full[name])
            for sid, c in ((0, gate), (1, up)):
                p = getattr(gu_l, name); p.weight_loader(p, c[name], sid)
        o_l.bias.weight_loader(o_l.bias, o_full["bias"])

        g = torch.Generator().manual_seed(3)
        x = o_full["x"]
        res = torch.randn(3, hidden, generator=g).to(dt)
        w_norm = (torch.rand(hidden, generator=g) + 0.5).to(dt)
        chunk = hidden // world
        # o_proj: all-reduce, THEN the residual; bias enters once (linear.py:1206-1213, llama.py:276-283)
        z, _, ssq = o_l(x[:, rank * chunk:(rank + 1) * chunk], residual=res, want_ssq=True)
        z_ref = (O.linear_fp32(x, o_full["codes"], o_full["codebooks"], o_full["scales"], [hidden], o_full["bias"])
                 .to(dt).float() + res.float())
        torch.testing.assert_close(z.float(), z_ref, rtol=2e-2, atol=2e-2)
        torch.testing.assert_close(ssq.sum(-1), z.float().pow(2).sum(-1), rtol=1e-4, atol=1e-4)
        # gate_up with RMSNorm from the producer's sum of squares == RMSNorm launch + plain linear
        y, _ = gu_l(z, input_norm=(ssq, w_norm, eps))
        xn, _ = O.fused_add_rms_norm(z, None, w_norm, eps)
        y_ref, _ = gu_l(xn)
        torch.testing.assert_close(y.float(), y_ref.float(), rtol=2e-2, atol=2e-2)
        # down_proj with SiLU*mul on its input, all-reduce, residual: against the full-tensor oracle
        out, _ = down_l(y, input_act="silu_and_mul", residual=z)
        acts = [torch.empty(3, inter // world, dtype=dt) for _ in range(world)]
        dist.all_gather(acts, O.silu_and_mul(y).contiguous())
        out_ref = O.linear_fp32(torch.cat(acts, -1), down_full["codes"], down_full["codebooks"], down_full["scales"],
                                [hidden], None).float() + z.float()
        torch.testing.assert_close(out.float(), out_ref, rtol=3e-2, atol=3e-2)
        same = [torch.empty_like(out) for _ in range(world)]
        dist.all_gather(same, out)
        assert torch.equal(same[0], same[1])

        # vocabulary-parallel embedding: 300 -> padded 320, 160 rows per rank; masked lookup + all-reduce
        emb_full = O.make_case([vocab], hidden, 8, 8, 1, dtype=dt, seed=15)
        emb = P.VocabParallelEmbedding(vocab, hidden, params_dtype=dt, quant_config=cfg)
        assert emb.num_embeddings_per_partition == 160 and tuple(emb.codes.shape) == (160, hidden // 8, 1)
        for name in ("codes", "codebooks", "scales"):
            p = getattr(emb, name); p.weight_loader(p, emb_full[name])
        ids = torch.tensor([[0, 159, 160, 299], [5, 200, 7, 161]])
        got = emb(ids)
        w_full = O.dequantize_partitions(emb_full["codes"], emb_full["codebooks"], [vocab], emb_full["scales"])
        assert torch.equal(got, w_full[ids])

        # set-up decisions are group-wide: one rank's failure turns the feature off everywhere
        from vllm_pq_b200 import distributed as D
        assert D._agree(True, dist.group.WORLD) is True
        assert D._agree(rank == 0, dist.group.WORLD) is False
        assert D.fused_allreduce_comm() is None and not D.oneshot_enabled()       # gloo: no peer-memory kernels

        # checkpoint walk at world size 2: this rank's vocabulary rows and experts are placed, the other
        # rank's experts are consumed without being reported (mixtral_quant.py:436-439)
        from tests.test_loader_cpu import TinyMoEModel, moe_checkpoint
        from vllm_pq_b200 import loader as L
        model = TinyMoEModel(P.PQConfig(8, 8, 1, 1, lm_head_quantized=True))
        tensors, full = moe_checkpoint((8, 8, 1))
        loaded, skipped = L.load_pq_linears(model, tensors.items())
        assert loaded == {n for n, _ in model.named_parameters()}
        assert skipped == [n for n in tensors if "input_layernorm" in n]
        mine = [2 * rank, 2 * rank + 1]
        moe = model.layers[1].block_sparse_moe
        assert moe.expert_indicies == mine and all((moe.experts[e] is not None) == (e in mine) for e in range(4))
        for e in mine:
            c = full[f"model.layers.1.block_sparse_moe.experts.{e}.w2"]
            assert torch.equal(moe.experts[e].w2.codes, c["codes"]) and torch.equal(moe.experts[e].w2.scales, c["scales"])
        c = full["lm_head"]                                   # 100 rows -> padded 128, 64 per rank
        have = 64 if rank == 0 else 36
        assert torch.equal(model.lm_head.codes[:have], c["codes"][64 * rank:64 * rank + have])
        assert not model.lm_head.scales[have:].any()
        ret[rank] = "ok"
    finally:
        P.destroy_tensor_parallel()
        dist.destroy_process_group()


def test_tp2_fused_chain_hooks_and_vocab_layers():
    world = 2
    ctx = mp.get_context("spawn")
    with ctx.Manager() as mgr:
        ret = mgr.dict()
        port = _free_port()
        procs = [ctx.Process(target=_worker_fused, args=(r, world, port, ret)) for r in range(world)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(timeout=300)
        assert dict(ret) == {0: "ok", 1: "ok"}, dict(ret)
