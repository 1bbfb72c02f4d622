set -x
python -m pytest tests -m gpu -x -q -k "mma or full_size or tc_sweep or golden_gemm or ragged" > gpurun_out/r1_03_pytest_mma.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/r1_03_pytest_mma.log
python tools/microbench.py --shapes qkv7b,o7b,gateup7b,down7b,gateup70b --formats 8x8x1,8x8x2 --tokens 1,2,4,8,16,32,64 --algos mma --pdl 1 --dequant 0 > gpurun_out/r1_03_micro_mma.jsonl 2>&1
for cfg in "1 1 8" "2 1 8" "4 1 8" "2 2 4" "2 4 2" "4 4 2" "2 8 1" "4 8 1"; do set -- $cfg
PQ_MMA_RT=$1 PQ_MMA_CS=$2 PQ_MMA_WK=$3 python tools/microbench.py --shapes o7b,gateup7b,down7b --formats 8x8x1 --tokens 1,8 --algos mma --pdl 1 --dequant 0 > gpurun_out/r1_03_micro_mma_rt$1_cs$2_wk$3.jsonl 2>&1
done
python tools/run_one.py --parts 11008,11008 --k 4096 --algo mma > gpurun_out/plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:pq_gemm_mma -s 4 -c 1 -o gpurun_out/r1_03_mma_gateup7b python tools/run_one.py --parts 11008,11008 --k 4096 --algo mma > gpurun_out/ncu2.log 2>&1
