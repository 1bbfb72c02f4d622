#!/usr/bin/env python
"""bench.py - Llama-2-7B PQ decode step (all PQ linears) on N GPUs of one node.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--batch B] [--format 8x8x1]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...
    python bench.py --impl reference        # the reference's torch PQ path on the host cores

A step = one pass of a batch of B tokens through the 4 PQ linears of all 32 layers
(qkv -> o -> gate_up -> down, TP = N: column-parallel QKV / gate-up, row-parallel O / down,
one all-reduce after each row-parallel layer).  Weights are synthetic (uniform random codes,
Gaussian codebooks) in the checkpoint layout; every layer has its own weights, so one step
streams the whole model (0.81 GB at 1 bit/weight) - far more than the 126 MB L2.

Printed JSON (rank 0): metric/value = tokens/s with inputs resident in HBM (CUDA-graph replay,
CUDA events, max over ranks); e2e = the same through host buffers (pinned x -> H2D -> step ->
D2H of the result, every step); roofline = the dominant kernel's algorithmic bytes / its
measured launch time against the measured HBM peak; cpu_baseline = the oracle's restatement of
the reference's generic_dequantize_gemm (same torch ops) on the host cores, bounded sample.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

METRIC = "llama2_7b_pq_decode_tokens_per_s"
UNIT = "tokens/s"


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.isfile(path):
        with open(path) as fh:
            p = json.load(fh)
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def measured_tc_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.isfile(path):
        with open(path) as fh:
            return float(json.load(fh)["bf16_tflops"])
    return 1650.0


def measured_traffic(model: str, fmt: str, dtype: str, world: int, batch: int, kernel: str):
    """DRAM bytes per launch of `kernel` from the committed ncu --set full capture, or None."""
    path = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    if not os.path.isfile(path):
        return None
    with open(path) as fh:
        table = json.load(fh)
    return table.get(f"{model}/{fmt}/{dtype}/tp{world}/batch{batch}", {}).get(kernel)


def parse_format(s: str):
    b, d, m = (int(v) for v in s.split("x"))
    return b, d, m


# ------------------------------------------------------------------------------------------
# clocks during the timed region
# ------------------------------------------------------------------------------------------
class ClockSampler:
    REASONS = {0x1: "gpu_idle", 0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown",
               0x10: "sync_boost", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
               0x80: "hw_power_brake_slowdown", 0x100: "display_clock_setting"}

    def __init__(self, index: int):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._h = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self._nv = pynvml
            self._h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self._h = None
        self._thr = threading.Thread(target=self._run, daemon=True)

    def _run(self):
        while not self._stop.is_set():
            try:
                self.samples.append(self._nv.nvmlDeviceGetClockInfo(self._h, self._nv.NVML_CLOCK_SM))
                mask = self._nv.nvmlDeviceGetCurrentClocksEventReasons(self._h)
                for bit, name in self.REASONS.items():
                    if mask & bit and name != "gpu_idle":
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.02)

    def __enter__(self):
        if self._h is not None:
            self._thr.start()
        return self

    def __exit__(self, *exc):
        self._stop.set()
        if self._h is not None:
            self._thr.join(timeout=1)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["unavailable"]}
        return {"sm_mhz": statistics.median(self.samples), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons)}


# ------------------------------------------------------------------------------------------
# CPU arm: the reference's torch PQ dequant+matmul on the host cores
# ------------------------------------------------------------------------------------------
def cpu_layer_time(fmt, batch, dtype, repeats, n_layers_total=32):
    """One decoder layer's four PQ linears through the oracle's literal restatement of
    generic_dequantize_gemm (aqlm.py:94-127: unpack -> embedding_bag -> scale -> F.linear).
    Returns seconds per layer (best of `repeats` after one warm-up)."""
    from oracle import pq_oracle as O
    b, d, m = fmt
    torch.set_num_threads(os.cpu_count() or 1)
    shapes = [([4096] * 3, 4096), ([4096], 4096), ([11008] * 2, 4096), ([4096], 11008)]
    cases = [O.make_case(parts, k, b, d, m, t=batch, dtype=dtype, seed=i) for i, (parts, k) in enumerate(shapes)]

    def run():
        for c in cases:
            O.generic_dequantize_gemm(c["x"], c["codes"], c["codebooks"], c["scales"],
                                      c["partition_sizes"], None, literal=True)
    run()
    times = []
    for _ in range(repeats):
        t0 = time.perf_counter()
        run()
        times.append(time.perf_counter() - t0)
    return min(times), statistics.median(times)


def run_reference_arm(args):
    """--impl reference: the reference's CPU implementation of the path on the host cores.  One step =
    one bounded sample of the workload: the four PQ linears of ONE decoder layer (1/32 of the
    Llama-2-7B step) through the oracle's literal port of generic_dequantize_gemm (aqlm.py:94-127:
    unpack, embedding_bag, scale, F.linear - /root/reference is absent on the GPU box).  Exactly
    --warmup untimed and --steps timed samples run; ms_per_step is what was timed, `value`
    extrapolates it to the whole 32-layer step (every layer does identical work)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import pq_oracle as O
    fmt = parse_format(args.format)
    b, d, m = fmt
    dtype = torch.float16 if args.dtype == "fp16" else torch.bfloat16
    layers = 32
    torch.set_num_threads(os.cpu_count() or 1)
    shapes = [([4096] * 3, 4096), ([4096], 4096), ([11008] * 2, 4096), ([4096], 11008)]
    cases = [O.make_case(parts, k, b, d, m, t=args.batch, dtype=dtype, seed=i) for i, (parts, k) in enumerate(shapes)]

    def sample():
        for c in cases:
            O.generic_dequantize_gemm(c["x"], c["codes"], c["codebooks"], c["scales"],
                                      c["partition_sizes"], None, literal=True)
    for _ in range(max(1, args.warmup)):
        sample()
    steps = max(1, min(args.steps, 200))
    t0 = time.perf_counter()
    for _ in range(steps):
        sample()
    sample_s = (time.perf_counter() - t0) / steps
    value = args.batch / (sample_s * layers)
    cores = torch.get_num_threads()
    desc = (f"one step = 1 of {layers} decoder layers (qkv,o,gate_up,down), {steps} timed; value extrapolates x{layers}; "
            f"oracle port of aqlm.py:94-127 (unpack, embedding_bag, scale, F.linear), {args.dtype}")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": steps, "warmup": max(1, args.warmup), "ms_per_step": sample_s * 1e3, "extrapolated_x": layers,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f16" if dtype == torch.float16 else "bf16", "data": "synthetic",
        # the same keys and workload description as the GPU arm prints (run_gpu_arm)
        "config": {"workload": f"{args.model} all PQ linears ({layers} layers x qkv,o,gate_up,down), decode",
                   "batch": args.batch, "format": f"{fmt[0]}bit x d{fmt[1]} x M{fmt[2]}",
                   "parallelism": f"tp{args.gpus}", "cold_l2": "n/a (host cores)", "launch": "n/a (host cores)",
                   "allreduce": "none", "mlp_act": "n/a (linears only)"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": desc},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


# ------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------
def capture(fn, stream):
    with torch.cuda.stream(stream):
        fn()
        fn()
        stream.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=stream):
            out = fn()
    return g, out


def time_replays(run, steps, world, device):
    """Exactly `steps` runs between events, barrier + synchronize on both sides, max over ranks."""
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        run()
    e1.record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], device=device, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return float(ms.item()) / steps


def check_result(model, x_dev, y_dev, world, n_layers=2):
    """Outside the timed region: the step's result is finite, and every linear of the first
    `n_layers` decoder layers - fed the input it actually saw in the stack - agrees with the
    oracle's fp64 evaluation of the same contraction within the P3 bound.  TP > 1:
    each rank checks its own shard's partial products (apply, before the all-reduce)."""
    from oracle import pq_oracle as O
    res = {"finite": bool(torch.isfinite(y_dev).all()), "linears": 0, "max_rms_rel": 0.0}
    tol = 2e-3 if y_dev.dtype == torch.float16 else 1.6e-2
    seen = []

    def hook(name):
        def fn(mod, inp, out):
            seen.append((name, mod, inp[0].detach().clone()))
        return fn
    handles = []
    for li, layer in enumerate(model.layers[:n_layers]):
        for name, lin in layer.linears():
            handles.append(lin.register_forward_hook(hook(f"{li}.{name}")))
    with torch.no_grad():
        model(x_dev)
        for name, lin, xin in seen:
            k = lin.codes.shape[1] * lin.codebooks.shape[3]
            xc = xin.cpu()
            if xin.shape[-1] == 2 * k:                      # --fused-act: down_proj was handed (gate | up)
                y = lin.quant_method.apply_silu_mul(lin, xin)
                xc = O.silu_and_mul(xc)
            else:
                y = lin.quant_method.apply(lin, xin)
            parts = list(getattr(lin.codebooks, "output_partition_sizes", []))
            ref = O.linear_fp32(xc, lin.codes.cpu(), lin.codebooks.cpu(), lin.scales.cpu(), parts, None)
            err = O.error_report(y.cpu(), ref)
            res["linears"] += 1
            res["max_rms_rel"] = max(res["max_rms_rel"], err["rms_rel"])
    for h in handles:
        h.remove()
    res["ok"] = res["finite"] and res["max_rms_rel"] < tol
    res["tol_rms_rel"] = tol
    return res


# ------------------------------------------------------------------------------------------
# the rest of BASELINE.json's metric, outside the headline's timed region (rank 0, one GPU)
# ------------------------------------------------------------------------------------------
def _time_sets(fn, sets, iters, stream):
    """One CUDA graph holding one call per (distinct) weight set - codes never stay L2-resident -
    replayed `iters` times between events on the launching stream; seconds per call."""
    with torch.cuda.stream(stream):
        for s_ in sets[:2]:
            fn(s_)
        stream.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=stream):
            keep = [fn(s_) for s_ in sets]
        for _ in range(3):
            g.replay()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        stream.synchronize()
        e0.record(stream)
        for _ in range(iters):
            g.replay()
        e1.record(stream)
        stream.synchronize()
    del keep
    return e0.elapsed_time(e1) * 1e-3 / (iters * len(sets))


def _make_sets(parts, k, b, d, m, dtype, device, min_bytes=256 << 20, max_sets=64):
    n, e = sum(parts), 1 << b
    code_bytes = n * (k // d) * m * (1 if b <= 8 else 2)
    nrep = max(2, min(max_sets, min_bytes // code_bytes + 1))
    sets = []
    for _ in range(nrep):
        codes = torch.randint(-(e // 2), e // 2, (n, k // d, m), device=device, dtype=torch.int32).to(
            torch.int8 if b <= 8 else torch.int16)
        cb = torch.randn(len(parts) * m, e, 1, d, device=device, dtype=dtype) / (k * m) ** 0.5
        sc = (torch.rand(n, 1, 1, 1, device=device) + 0.5).to(dtype)
        sets.append((codes, cb, sc))
    return sets, code_bytes


def _algo_bytes(t, n, k, d, b, m, p):
    return n * (k // d) * m * (1 if b <= 8 else 2) + p * m * (1 << b) * d * 2 + n * 2 + t * k * 2 + t * n * 2


def run_extras(P, S, device, stream, hbm_gbs, tc_peak, clk_index):
    """configs 1, 3, 5, the 70B layer shapes and the same-box reference kernels."""
    out = {}
    dt = torch.float16

    def gemv(parts, k, b, d, m, t=1, iters=20):
        sets, _ = _make_sets(parts, k, b, d, m, dt, device)
        x = torch.randn(t, k, device=device, dtype=dt)
        sec = _time_sets(lambda s_: P.ops.pq_gemm(x, s_[0], s_[1], s_[2], parts, None), sets, iters, stream)
        ab = _algo_bytes(t, sum(parts), k, d, b, m, len(parts))
        row = {"us": sec * 1e6, "gbs": ab / sec / 1e9, "frac": ab / sec / 1e9 / hbm_gbs}
        return row, sets, x

    # config 1: single 4096 x 4096 linear, 8 bit x d8 x M1, one token
    out["gemv_single"], _, _ = gemv([4096], 4096, 8, 8, 1)
    out["gemv_single"]["workload"] = "config 1: 4096x4096, 8bit x d8 x M1, batch 1"
    # config 5: 4/8/12-bit codes x subvector 4/8/16, 4096 x 4096, one token
    out["sweep"] = {}
    for b in (4, 8, 12):
        for d in (4, 8, 16):
            out["sweep"][f"{b}x{d}x1"] = gemv([4096], 4096, b, d, 1, iters=10)[0]
    # Llama-2-70B layer shapes, one token, 8 bit x d8 x M1
    out["llama70b_layer"] = {}
    for name, parts, k in (("qkv_proj", [8192, 1024, 1024], 8192), ("o_proj", [8192], 8192),
                           ("gate_up_proj", [28672] * 2, 8192), ("down_proj", [8192], 28672)):
        out["llama70b_layer"][name] = gemv(parts, k, 8, 8, 1, iters=10)[0]
    # config 3: Llama-3-8B MLP shapes, 2048 tokens (tcgen05 GEMM), with the same-box baselines
    have_ref = False
    try:
        from oracle import ref_cuda as R                    # checker / baseline only
        have_ref = R.available()
    except Exception:
        R = None
    out["prefill"] = {}
    with ClockSampler(clk_index) as clk:
        for name, parts, k in (("gate_up8b", [14336] * 2, 4096), ("down8b", [4096], 14336)):
            n, t = sum(parts), 2048
            for (b, d, m) in ((8, 8, 1), (8, 8, 2), (16, 8, 1)):
                sets, _ = _make_sets(parts, k, b, d, m, dt, device, max_sets=8)
                x = torch.randn(t, k, device=device, dtype=dt)
                sec = _time_sets(lambda s_: P.ops.pq_gemm(x, s_[0], s_[1], s_[2], parts, None), sets, 5, stream)
                row = {"us": sec * 1e6, "tflops": 2 * t * n * k / sec / 1e12, "frac": 2 * t * n * k / sec / 1e12 / tc_peak}
                if have_ref and R.supports(b, d, m, dt):
                    rs = _time_sets(lambda s_: R.optimized_dequantize_gemm(x, s_[0], s_[1], s_[2], parts, None),
                                    sets[:3], 3, stream)
                    row["ref_cuda_dequant_cublas_us"] = rs * 1e6
                if (b, d, m) == (8, 8, 1):
                    dense = [(P.ops.pq_dequant(s_[0], s_[1], parts, s_[2]),) for s_ in sets[:3]]
                    ds = _time_sets(lambda s_: torch.matmul(x, s_[0].t()), dense, 5, stream)
                    row["dense_cublas_us"] = ds * 1e6
                    del dense
                out["prefill"][f"{name}/{b}x{d}x{m}"] = row
                del sets
    out["prefill"]["clocks"] = clk.summary()
    out["prefill"]["peak_tflops"] = tc_peak
    # same-box reference kernels at decode (aqlm_gemm: one token, the two formats it supports)
    if have_ref:
        out["ref_cuda"] = {}
        for name, parts, k in (("qkv7b", [4096] * 3, 4096), ("down7b", [4096], 11008), ("gate_up70b", [28672] * 2, 8192)):
            for (b, d, m) in ((8, 8, 2), (16, 8, 1)):
                ours, sets, x = gemv(parts, k, b, d, m, iters=10)
                rs = _time_sets(lambda s_: R.aqlm_gemm(x, s_[0], s_[1], s_[2], parts, None), sets, 5, stream)
                out["ref_cuda"][f"{name}/{b}x{d}x{m}/t1"] = {"ours_us": ours["us"], "aqlm_gemm_us": rs * 1e6}
                del sets
    torch.cuda.empty_cache()
    return out


def run_gpu_arm(args):
    import vllm_pq_b200 as P
    from vllm_pq_b200 import stack as S

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=device)
        P.init_tensor_parallel()
    assert world == args.gpus, f"--gpus {args.gpus} but WORLD_SIZE={world}"

    fmt = parse_format(args.format)
    b, d, m = fmt
    dtype = torch.float16 if args.dtype == "fp16" else torch.bfloat16
    P.ops.set_pdl(bool(args.pdl))
    shape = S.SHAPES[args.model]
    cfg = P.PQConfig(d, b, m, 1)
    with torch.device(device):
        model = S.PQLlamaLinearStack(shape, cfg, dtype, fused_act=bool(args.fused_act)).init_synthetic(seed=rank)
    B = args.batch
    stream = torch.cuda.Stream()

    x_host = torch.randn(B, shape.hidden, dtype=dtype).pin_memory()
    y_host = torch.empty(B, shape.hidden, dtype=dtype).pin_memory()
    x_dev = x_host.to(device)

    # torch's NCCL process group does not replay reliably from a captured CUDA graph (observed: hang at
    # TP=2 with PQ_ONESHOT_ALLREDUCE=0); the one-shot NVLink kernel does.  Without it, TP > 1 runs eagerly.
    use_graph = world == 1 or P.distributed.oneshot_enabled()
    with torch.no_grad():
        if use_graph:
            graph, y_dev = capture(lambda: model(x_dev), stream)
        else:
            graph = None
            with torch.cuda.stream(stream):
                y_dev = model(x_dev)

        # per-projection graphs: the same kernel over all layers' weights (cold L2), for the roofline
        kernel_rows = []
        if rank == 0 or world > 1:
            for name in ("qkv_proj", "o_proj", "gate_up_proj", "down_proj"):
                lins = [getattr(layer, name) for layer in model.layers]
                k_in = lins[0].codes.shape[1] * d
                xin = torch.randn(B, k_in, dtype=dtype, device=device)
                g_k, outs = capture(lambda xin=xin, lins=lins: [lin.quant_method.apply(lin, xin) for lin in lins], stream)
                # the graph holds raw pointers: its input and outputs must outlive it
                kernel_rows.append((name, lins[0], g_k, len(lins), (xin, outs)))

    def run_step():
        if graph is not None:
            graph.replay()
            return y_dev
        with torch.no_grad():
            return model(x_dev)

    def step_resident():
        with torch.cuda.stream(stream):
            run_step()

    def step_e2e():
        with torch.cuda.stream(stream):
            x_dev.copy_(x_host, non_blocking=True)
            y_host.copy_(run_step(), non_blocking=True)

    def timed(run, steps):
        # events must be recorded on the launching stream
        with torch.cuda.stream(stream):
            return time_replays(run, steps, world, device)

    for _ in range(max(args.warmup, 3)):
        step_e2e()
    # let clocks settle on a fresh box: a FIXED number of replays - a step contains collectives at
    # TP > 1, so every rank must run exactly the same number of them (a time-based loop deadlocks)
    for _ in range(300):
        step_resident()
    torch.cuda.synchronize()

    with ClockSampler(local) as clk:
        ms_step = timed(step_resident, args.steps)
        ms_e2e = timed(step_e2e, args.steps)
        kernels = {}
        hbm_gbs, peak_src = measured_peaks()
        for name, lin, g_k, n_l, _keep in kernel_rows:
            ms_k = timed(lambda: g_k.replay(), max(10, args.steps // 4)) / n_l
            ab = S.algorithmic_bytes(lin, B)
            kernels[name] = {"us": ms_k * 1e3, "algorithmic_bytes": ab, "gbs": ab / (ms_k * 1e-3) / 1e9}
    clocks = clk.summary()
    checked = check_result(model, x_dev, y_dev, world)      # outside the timed region
    extras = {}
    if world == 1 and not args.no_extras:
        # decode at batch 16 and 64 through the same stack (config 2's other batch sizes)
        extras["batch"] = {}
        for bt in (16, 64):
            xb = torch.randn(bt, shape.hidden, dtype=dtype, device=device)
            with torch.no_grad():
                gb, _ = capture(lambda: model(xb), stream)
            ms_b = timed(lambda: gb.replay(), max(5, args.steps // 5))
            extras["batch"][str(bt)] = {"tokens_per_s": bt / (ms_b * 1e-3), "ms_per_step": ms_b}
            del gb
        extras.update(run_extras(P, S, device, stream, hbm_gbs, measured_tc_peak(), local))
    if world > 1:                                           # every rank must hold the same bits
        ref_y = y_dev.clone()
        dist.broadcast(ref_y, src=0)
        same = torch.tensor([int(torch.equal(ref_y, y_dev) and checked["ok"])], device=device)
        dist.all_reduce(same, op=dist.ReduceOp.MIN)
        checked["identical_across_ranks_and_ok_on_all"] = bool(same.item())
        checked["ok"] = checked["ok"] and bool(same.item())

    if rank == 0:
        dom = max(kernels, key=lambda k: kernels[k]["us"])
        tot_us = sum(v["us"] for v in kernels.values())
        for v in kernels.values():
            v["share_of_step"] = v["us"] / tot_us
        achieved = kernels[dom]["gbs"]
        out = {
            "metric": METRIC, "value": B / (ms_step * 1e-3), "unit": UNIT, "n_gpus": world,
            "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_step,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f16" if dtype == torch.float16 else "bf16",
            "data": "synthetic",
            "config": {"workload": f"{shape.name} all PQ linears ({len(model.layers)} layers x qkv,o,gate_up,down), decode",
                       "batch": B, "format": f"{b}bit x d{d} x M{m}", "parallelism": f"tp{world}",
                       "cold_l2": f"each step streams {model.code_bytes() * world / 1e6:.0f} MB of distinct codes (> 126 MB L2)",
                       "launch": ("CUDA graph" if use_graph else "eager") + (" + programmatic dependent launch" if args.pdl else ""),
                       "allreduce": ("none" if world == 1 else "nccl" if not P.distributed.oneshot_enabled() else
                                     "fused into the row-parallel GEMV epilogue (NVLink peer memory, LL packets)"
                                     if (B <= 8 and b == 8 and P.distributed.fused_allreduce_comm() is not None) else
                                     "one-shot NVLink P2P kernel" if B * shape.hidden * 2 <= (1 << 20) else
                                     "two-stage NVLink P2P kernel (reduce-scatter + all-gather)"),
                       "mlp_act": ("SiLU(gate)*up on down_proj's x load" if args.fused_act
                                   else "stand-in slice of gate_up (no activation math)")},
            "e2e": {"value": B / (ms_e2e * 1e-3), "unit": UNIT,
                    "h2d_bytes_per_step": x_host.numel() * x_host.element_size(),
                    "d2h_bytes_per_step": y_host.numel() * y_host.element_size()},
            "gpu_launches": model.launches_per_step() * args.steps,
            "checked": checked["ok"], "check": checked,
            "clocks": clocks,
            "roofline": {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": hbm_gbs, "unit": "GB/s",
                         "frac": achieved / hbm_gbs,
                         "traffic": measured_traffic(args.model, args.format, args.dtype, world, B, dom),
                         "algorithmic_bytes": kernels[dom]["algorithmic_bytes"], "peak_source": peak_src,
                         "whole_step_gbs": sum(v["algorithmic_bytes"] for v in kernels.values()) * len(model.layers)
                         / (ms_step * 1e-3) / 1e9},
            "kernels": kernels,
        }
        out.update(extras)
        if world == 1 and not args.no_cpu:
            best, med = cpu_layer_time(fmt, B, dtype, repeats=3)
            out["cpu_baseline"] = {
                "value": B / (med * shape.layers), "unit": UNIT, "cores": torch.get_num_threads(), "kind": "port",
                "sample": f"1 of {shape.layers} decoder layers timed (best {best * 1e3:.0f} ms, median {med * 1e3:.0f} ms), "
                          f"x{shape.layers}; oracle port of aqlm.py:94-127 with the reference's torch ops"}
        print(json.dumps(out))
    if world > 1:
        P.destroy_tensor_parallel()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=1)
    ap.add_argument("--format", default="8x8x1", help="nbits x d x M; 8x8x1 is BASELINE config 1's format")
    ap.add_argument("--dtype", default="fp16", choices=["fp16", "bf16"])
    ap.add_argument("--model", default="llama-2-7b")
    ap.add_argument("--pdl", type=int, default=1)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-extras", action="store_true",
                    help="skip configs 1/3/5, batch 16/64, the 70B layer shapes and the same-box reference kernels")
    ap.add_argument("--fused-act", type=int, default=0,
                    help="1: down_proj takes SiLU(gate)*up on its x load (SURVEY 8f N2) instead of the slicing stand-in")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_gpu_arm(args)


if __name__ == "__main__":
    main()
