set -x
python -m pytest tests -m gpu -x -q > gpurun_out/r1_04_pytest.log 2>&1; echo "pytest rc=$?"
tail -4 gpurun_out/r1_04_pytest.log
python bench.py --steps 30 --warmup 5 > gpurun_out/r1_04_bench.json 2> gpurun_out/r1_04_bench.err; echo "bench rc=$?"
python tools/microbench.py --shapes qkv7b,o7b,gateup7b,down7b,gateup70b,down70b --formats 8x8x1,8x8x2,16x8x1 --tokens 1,4,16,64 --pdl 1 --dequant 0 --ref 1 > gpurun_out/r1_04_micro_decode.jsonl 2>&1
for cfg in "1 1" "2 1" "4 1" "8 1" "2 2" "4 2" "8 2"; do set -- $cfg
PQ_T1_ALGO=2 PQ_LUT_CS=$1 PQ_LUT_OCC=$2 python tools/microbench.py --shapes qkv7b,o7b,gateup7b,down7b --formats 8x8x1 --tokens 1 --algos lut --pdl 1 --dequant 0 > gpurun_out/r1_04_lut_cs$1_occ$2.jsonl 2>&1
done
