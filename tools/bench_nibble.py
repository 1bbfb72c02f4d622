"""4-bit codes at decode: the checkpoint layout (one int8 per 4-bit code) against the nibble-packed private
copy + product codebook built by process_weights_after_loading.  One token, rotating weight sets, graph."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import vllm_pq_b200 as P  # noqa: E402


def time_graph(fn, n_sets, iters=20):
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        for i in range(min(2, n_sets)):
            fn(i)
        st.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=st):
            keep = [fn(i) for i in range(n_sets)]
        for _ in range(3):
            g.replay()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        st.synchronize()
        e0.record(st)
        for _ in range(iters):
            g.replay()
        e1.record(st)
        st.synchronize()
    del keep
    return e0.elapsed_time(e1) * 1e3 / (iters * n_sets)


def main():
    out = []
    for name, parts, k in (("gate_up70b", [28672] * 2, 8192), ("gate_up7b", [11008] * 2, 4096)):
        for (d, m) in ((8, 1), (8, 2)):
            cfg = P.PQConfig(d, 4, m, 1)
            n = sum(parts)
            n_sets = max(2, min(16, (256 << 20) // (n * (k // d) * m) + 1))
            layers = []
            for _ in range(n_sets):
                with torch.device("cuda"):
                    lin = P.MergedColumnParallelLinear(k, parts, bias=False, params_dtype=torch.float16, quant_config=cfg)
                lin.codes.data.copy_(torch.randint(-8, 8, lin.codes.shape, device="cuda", dtype=torch.int32))
                lin.codebooks.data.copy_(torch.randn(lin.codebooks.shape, device="cuda") / (k * m) ** 0.5)
                lin.scales.data.copy_(torch.rand(lin.scales.shape, device="cuda") + 0.5)
                layers.append(lin)
            x = torch.randn(1, k, device="cuda", dtype=torch.float16)
            us_plain = time_graph(lambda i: layers[i](x)[0], n_sets)
            for lin in layers:
                lin.quant_method.process_weights_after_loading(lin)
            us_packed = time_graph(lambda i: layers[i](x)[0], n_sets)
            weights = n * k
            out.append({"layer": name, "format": f"4x{d}x{m}", "unpacked_us": us_plain, "packed_us": us_packed,
                        "speedup": us_plain / us_packed, "gweights_per_s_packed": weights / us_packed / 1e3,
                        "packed_code_gbs": layers[0].codes_nibble.numel() / us_packed / 1e3})
            print(json.dumps(out[-1]), flush=True)
            del layers
            torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
