"""Multi-GPU check of the tensor-parallel PQ path; run under torchrun, one rank per GPU:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29511 tools/tp_check.py

1. one-shot NVLink all-reduce (csrc/pq_allreduce.cu) == NCCL all-reduce, bit for bit across ranks,
   eager and under CUDA-graph replay, for decode-sized messages;
2. a TP = N decoder-layer stack (column-parallel qkv / gate_up, row-parallel o / down + reduce),
   loaded through the weight loaders from FULL tensors, reproduces the TP = 1 result computed on
   rank 0's GPU from the same tensors, within the fp16 tolerance of the parity tests.
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
    for n in (4096, 8192, 64 * 4096):
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
    # graph replay of the one-shot kernel (epoch counter lives on the device)
    x = torch.randn(8192, device=dev, dtype=torch.float16)
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
        print(f"[rank {rank}] graph-replayed all-reduce mismatch")

    # ---- 2. TP stack vs TP = 1 ----
    for (b, d, m) in ((8, 8, 1), (8, 8, 2), (16, 8, 1)):
        hidden, head, heads, kv_heads, inter = 1024, 64, 16, 8, 2048
        dt = torch.float16
        cfg = P.PQConfig(d, b, m, 1)
        kw = dict(bias=False, params_dtype=dt, quant_config=cfg)
        full = {name: O.make_case(parts, k, b, d, m, t=4, dtype=dt, seed=sd)
                for sd, (name, parts, k) in enumerate((("q", [heads * head], hidden), ("k", [kv_heads * head], hidden),
                                                       ("v", [kv_heads * head], hidden), ("o", [hidden], heads * head),
                                                       ("gate", [inter], hidden), ("up", [inter], hidden),
                                                       ("down", [hidden], inter)))}
        for c in full.values():                   # keep activations O(1) through the four linears (fp16 range)
            kdim = c["codes"].shape[1] * d
            c["codebooks"] = (c["codebooks"].float() / (kdim * m) ** 0.5).to(dt)
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
        x = full["q"]["x"].to(dev)
        # TP path: qkv (column) -> o on the q slice (row + reduce) -> gate_up (column) -> down on gate slice (row)
        y_qkv, _ = qkv(x)
        qw = heads * head // world
        h, _ = o(y_qkv[:, :qw].contiguous())
        y_gu, _ = gu(h)
        iw = inter // world
        out, _ = down(y_gu[:, :iw].contiguous())
        torch.cuda.synchronize()

        # TP = 1 reference on this GPU from the same full tensors, through the same kernels
        def lin(key, xin):
            c = full[key]
            return P.ops.pq_gemm(xin, c["codes"].to(dev), c["codebooks"].to(dev), c["scales"].to(dev),
                                 c["partition_sizes"], None)
        h1 = lin("o", lin("q", x))
        out1 = lin("down", lin("gate", h1))
        torch.cuda.synchronize()
        scale = float(out1.float().abs().max())
        err = float((out.float() - out1.float()).abs().max()) / scale
        if not (err <= 5e-3):                       # also catches NaN
            ok = False
            print(f"[rank {rank}] TP stack {b}x{d}x{m}: rel err {err:.2e}")
        elif rank == 0:
            print(f"tp{world} stack {b}bit d{d} M{m}: max rel err vs tp1 = {err:.2e}")

    flag = torch.tensor([1 if ok else 0], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if rank == 0:
        print(f"TP_CHECK {'PASS' if int(flag.item()) else 'FAIL'} world={world} oneshot={oneshot}")
    P.destroy_tensor_parallel()
    dist.destroy_process_group()
    sys.exit(0 if int(flag.item()) else 1)


if __name__ == "__main__":
    main()
