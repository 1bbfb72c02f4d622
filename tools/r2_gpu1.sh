# GPU call 1 of round 2: full GPU suite, smoke, bench, same-box reference matrices (decode + prefill)
set -x
mkdir -p gpurun_out
O=gpurun_out/r2_10
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm,power.limit --format=csv | tee ${O}_gpu.txt
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -6 | tee ${O}_pytest_tail.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -8 | tee ${O}_smoke.txt
timeout 600 python bench.py > ${O}_bench.json 2> ${O}_bench.err; echo "bench rc=$?"; cut -c1-600 ${O}_bench.json
timeout 300 python bench.py --no-cpu --fused-act 1 > ${O}_bench_fused_act.json 2>> ${O}_bench.err; cut -c1-300 ${O}_bench_fused_act.json
# prefill (config 3): ours vs reference dequant + cuBLAS vs dense cuBLAS
timeout 600 python tools/microbench.py --shapes gateup8b,down8b --formats 8x8x1,8x8x2,16x8x1 --tokens 2048 \
   --dequant 0 --ref 1 --dense 1 --dtype fp16 --out ${O}_microbench_prefill_vs_ref.json 2>&1 | tail -30
# decode matrix vs the reference kernels
timeout 900 python tools/microbench.py --shapes qkv7b,o7b,gateup7b,down7b,gateup70b,down70b --formats 8x8x2,16x8x1 \
   --tokens 1,2,4,8,16,32,64 --dequant 0 --ref 1 --warm 0.15 --out ${O}_microbench_decode_vs_ref.json 2>&1 | tail -100
