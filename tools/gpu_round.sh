set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -k "lut_tma" > gpurun_out/r1_06_pytest_lut2.log 2>&1; echo "pytest rc=$?"
tail -15 gpurun_out/r1_06_pytest_lut2.log
for shp in "22016 4096 2" "4096 11008 1" "4096 4096 1" "12288 4096 3"; do
  set -- $shp
  timeout 120 python tools/trace_lut.py --n $1 --k $2 --parts $3 --algo 5 >> gpurun_out/r1_06_trace_lut2.txt 2>&1
done
tail -40 gpurun_out/r1_06_trace_lut2.txt
timeout 600 python tools/microbench.py --shapes qkv7b,o7b,gateup7b,down7b,gateup70b,down70b --formats 8x8x1,8x8x2 --tokens 1,2,4 --algos auto,lut2 --pdl 1 --dequant 0 > gpurun_out/r1_06_microbench_lut2.jsonl 2>&1; echo "microbench rc=$?"
grep -c gemm gpurun_out/r1_06_microbench_lut2.jsonl
PQ_T1_ALGO=5 timeout 300 python bench.py --steps 20 --warmup 5 > gpurun_out/r1_06_bench_lut2.json 2> gpurun_out/r1_06_bench_lut2.err; echo "bench lut2 rc=$?"
cat gpurun_out/r1_06_bench_lut2.json | head -c 1500
