"""Multi-GPU check of the tensor-parallel PQ path; run under torchrun, one rank per GPU:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29511 tools/tp_check.py

1. one-shot NVLink all-reduce (csrc/pq_allreduce.cu) == NCCL all-reduce, bit for bit across ranks,
   eager and under CUDA-graph replay, for decode-sized messages;
2. the TP = N linears of a decoder layer (column-parallel qkv / gate_up, row-parallel o / down),
   loaded through the weight loaders from FULL tensors, against the ORACLE's fp64 contraction over
   the full tensors: plain, with the all-reduce fused into the GEMV epilogue (+ residual add + sum
   of squares), RMSNorm and SiLU*mul on the x loads; bitwise identical across ranks; fused ==
   separate launches within fp16 rounding; CUDA-graph replay of the fused chain == eager.
Exit code 0 and one "TP_CHECK PASS" line from rank 0 on success.
"""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import vllm_pq_b200 as P  # noqa: E402
from oracle import pq_oracle as O  # noqa: E402  (checker only)


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    P.init_tensor_parallel()
    ok = True

    # ---- 1. one-shot all-reduce vs NCCL ----
    oneshot = P.distributed.oneshot_enabled()
    # one-shot (<= 1 MiB), two-stage (<= 64 MiB: 2048 x 4096, 2048 x 8192, a size that does not divide
    # into equal chunks) and NCCL (above)
    for n in (4096, 8192, 64 * 4096, 2048 * 4096, 2048 * 8192, 1000 * 4096 + 8, (64 << 19) + 8):
        for dt in (torch.float16, torch.bfloat16):
            g = torch.Generator(device=dev).manual_seed(100 + rank)
            x = torch.randn(n, device=dev, dtype=dt, generator=g)
            ref = x.float().clone()
            dist.all_reduce(ref)
            for _ in range(3):                                   # epochs alternate staging slots
                got = P.tensor_model_parallel_all_reduce(x.clone())
            torch.cuda.synchronize()
            err = float((got.float() - ref).abs().max() / ref.abs().max())
            tol = 2e-3 if dt == torch.float16 else 1.6e-2
            same = [torch.empty_like(got) for _ in range(world)]
            dist.all_gather(same, got)
            bitwise = all(torch.equal(same[0], s) for s in same)
            if not (err <= tol) or not bitwise:
                ok = False
                print(f"[rank {rank}] allreduce n={n} {dt}: err={err:.2e} bitwise_equal_across_ranks={bitwise}")
    # graph replay of the one-shot and of the two-stage kernel (epoch counters live on the device)
    for n_graph in (8192, 2048 * 8192):
        x = torch.randn(n_graph, device=dev, dtype=torch.float16)
        s = torch.cuda.Stream()
        with torch.cuda.stream(s):
            P.tensor_model_parallel_all_reduce(x.clone())
            s.synchronize()
            dist.barrier()
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr, stream=s):
                y = P.tensor_model_parallel_all_reduce(x)
        ref = x.float().clone()
        dist.all_reduce(ref)
        for _ in range(5):
            gr.replay()
        torch.cuda.synchronize()
        if not (float((y.float() - ref).abs().max() / ref.abs().max()) <= 2e-3):
            ok = False
            print(f"[rank {rank}] graph-replayed all-reduce mismatch (n = {n_graph})")

    # ---- 2. TP decoder-layer linears against the ORACLE (fp64 contraction over the FULL tensors) ----
    # hidden 2048, intermediate 5632: at TP = 8 down_proj's shard has 88 code bytes per row (not a
    # multiple of 16: the cp.async loader), o_proj's 32 (a single partial tile).
    EPS = 1e-5
    hidden, head, heads, kv_heads, inter = 2048, 128, 16, 8, 5632
    dt = torch.float16
    tol = 2e-3

    def gather_cols(x):                               # [t, w] per rank -> [t, world * w]
        parts = [torch.empty_like(x) for _ in range(world)]
        dist.all_gather(parts, x.contiguous())
        return torch.cat(parts, -1)

    def same_on_all_ranks(x):
        parts = [torch.empty_like(x) for _ in range(world)]
        dist.all_gather(parts, x.contiguous())
        return all(torch.equal(parts[0], q) for q in parts)

    def report(name, got, ref64, extra=""):
        nonlocal ok
        err = O.error_report(got.cpu(), ref64)
        good = err["rms_rel"] < tol
        if not good:
            ok = False
        if not good or rank == 0:
            print(f"[rank {rank}] tp{world} {name}{extra}: rms_rel vs oracle {err['rms_rel']:.2e} {'ok' if good else 'FAIL'}")

    for (b, d, m) in ((8, 8, 1), (8, 8, 2), (16, 8, 1)):
        cfg = P.PQConfig(d, b, m, 1)
        kw = dict(bias=False, params_dtype=dt, quant_config=cfg)
        full = {name: O.make_case(parts, k, b, d, m, t=16, dtype=dt, seed=sd)
                for sd, (name, parts, k) in enumerate((("q", [heads * head], hidden), ("k", [kv_heads * head], hidden),
                                                       ("v", [kv_heads * head], hidden), ("o", [hidden], heads * head),
                                                       ("gate", [inter], hidden), ("up", [inter], hidden),
                                                       ("down", [hidden], inter)))}
        for c in full.values():                   # zero-mean codebooks, activations O(1) (stack.py init_synthetic)
            kdim = c["codes"].shape[1] * d
            cb = c["codebooks"].float() / (kdim * m) ** 0.5
            c["codebooks"] = (cb - cb.mean(1, keepdim=True)).to(dt)
        with torch.device(dev):
            qkv = P.QKVParallelLinear(hidden, head, heads, kv_heads, **kw)
            o = P.RowParallelLinear(heads * head, hidden, **kw)
            gu = P.MergedColumnParallelLinear(hidden, [inter, inter], **kw)
            down = P.RowParallelLinear(inter, hidden, **kw)
        for name in ("codes", "codebooks", "scales"):
            for sid in ("q", "k", "v"):
                p = getattr(qkv, name); p.weight_loader(p, full[sid][name].to(dev), sid)
            for sid, key in ((0, "gate"), (1, "up")):
                p = getattr(gu, name); p.weight_loader(p, full[key][name].to(dev), sid)
            for layer, key in ((o, "o"), (down, "down")):
                p = getattr(layer, name); p.weight_loader(p, full[key][name].to(dev))
        lin64 = lambda key, xin: O.linear_fp32(xin.cpu(), full[key]["codes"], full[key]["codebooks"],   # noqa: E731
                                               full[key]["scales"], full[key]["partition_sizes"], None)
        norm_w = (torch.rand(hidden, generator=torch.Generator().manual_seed(7)) + 0.5).to(dt)
        norm_w_dev = norm_w.to(dev)
        qw, kvw, iw = heads * head // world, max(1, kv_heads // world) * head, inter // world
        for t in (1, 4, 16):
            x = full["q"]["x"][:t].to(dev)
            res = full["o"]["x"][:t, :hidden].to(dev)
            tag = f" {b}x{d}x{m} t={t}"
            # column-parallel qkv: this rank's q rows (kv heads are divided or replicated, checked in the gloo tests)
            y_qkv, _ = qkv(x)
            report("qkv(q)", y_qkv[:, :qw], lin64("q", x)[:, rank * qw:(rank + 1) * qw], tag)
            # row-parallel o_proj + all-reduce (fused in the GEMV epilogue for t <= 8, nbits <= 8)
            attn = y_qkv[:, :qw].contiguous()
            h, _ = o(attn)
            h_ref = lin64("o", gather_cols(attn))
            report("o_proj+allreduce", h, h_ref, tag)
            if not same_on_all_ranks(h):
                ok = False
                print(f"[rank {rank}] o_proj result differs across ranks{tag}")
            # the same with the residual add and the sum of squares in the epilogue, RMSNorm on gate_up's x load
            z, _, ssq = o(attn, residual=res, want_ssq=True)
            report("o_proj+allreduce+residual", z, h_ref + res.double().cpu(), tag)
            y_gu, _ = gu(z, input_norm=(ssq, norm_w_dev, EPS))
            xn_ref, _ = O.fused_add_rms_norm(z.cpu(), None, norm_w, EPS)
            g_ref = torch.cat([lin64("gate", xn_ref)[:, rank * iw:(rank + 1) * iw],
                               lin64("up", xn_ref)[:, rank * iw:(rank + 1) * iw]], -1)
            report("rmsnorm+gate_up", y_gu, g_ref, tag)
            # row-parallel down_proj with SiLU(gate) * up on its x load, + all-reduce + residual
            out, _ = down(y_gu, input_act="silu_and_mul", residual=z)
            act = O.silu_and_mul(y_gu.cpu())
            out_ref = lin64("down", gather_cols(act.to(dev))) + z.double().cpu()
            report("silu_mul+down_proj+allreduce+residual", out, out_ref, tag)
            if not same_on_all_ranks(out):
                ok = False
                print(f"[rank {rank}] down_proj result differs across ranks{tag}")
            # fused epilogue == separate launches (GEMV, one-shot all-reduce, add) up to fp16 rounding of the partials
            os.environ["PQ_FUSED_ALLREDUCE"] = "0"
            h_sep, _ = o(attn)
            os.environ["PQ_FUSED_ALLREDUCE"] = "1"
            rel = float((h.double() - h_sep.double()).pow(2).mean().sqrt() / h_sep.double().pow(2).mean().sqrt())
            if not rel < 1e-3:
                ok = False
                print(f"[rank {rank}] fused vs separate all-reduce{tag}: {rel:.2e}")
        # CUDA-graph replay of the fused chain (device-side epoch): 6 replays == eager
        x1 = full["q"]["x"][:1].to(dev)
        res1 = full["o"]["x"][:1, :hidden].to(dev)

        def chain():
            yq, _ = qkv(x1)
            zz, _, sq = o(yq[:, :qw].contiguous(), residual=res1, want_ssq=True)
            yg, _ = gu(zz, input_norm=(sq, norm_w_dev, EPS))
            oo, _ = down(yg, input_act="silu_and_mul", residual=zz)
            return oo
        st = torch.cuda.Stream()
        with torch.cuda.stream(st):
            eager = chain().clone()
            chain()
            st.synchronize()
            dist.barrier()
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr, stream=st):
                outg = chain()
            for _ in range(6):
                gr.replay()
            st.synchronize()
        if not torch.equal(outg, eager):
            ok = False
            print(f"[rank {rank}] graph replay of the fused chain != eager ({b}x{d}x{m})")
        dist.barrier()

    flag = torch.tensor([1 if ok else 0], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        print(f"TP_CHECK {'PASS' if int(flag.item()) else 'FAIL'} world={world} oneshot={oneshot}")
    P.destroy_tensor_parallel()
    dist.destroy_process_group()
    sys.exit(0 if int(flag.item()) else 1)


if __name__ == "__main__":
    main()
