mkdir -p gpurun_out
O=gpurun_out/r2_28
timeout 300 python tools/bench_norm.py | tee ${O}_bench_norm.json

timeout 300 python tools/bench_norm.py --shape llama-2-70b --layers 16 | tee -a ${O}_bench_norm.json
