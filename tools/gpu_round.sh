set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_ref_cuda.py -x -q -k "test_gemv_matches_reference_cuda" 2>&1 | tail -4
PQ_AUTO_LUT2=0 timeout 300 python -m pytest tests/test_gpu_ref_cuda.py -x -q -k "test_gemv_matches_reference_cuda" 2>&1 | tail -4
timeout 300 python -m pytest "tests/test_gpu_ref_cuda.py::test_gemv_matches_reference_cuda[False-parts1-2048-1-fmt1]" -x -q 2>&1 | tail -4
timeout 600 compute-sanitizer --tool memcheck python tools/run_one.py --parts 1024,512,512 --k 2048 --fmt 8x8x2 --algo lut2 --nsets 2 --iters 2 > gpurun_out/r1_17_sanitizer.log 2>&1; echo "sanitizer rc=$?"
tail -15 gpurun_out/r1_17_sanitizer.log
for cs in 4 8; do
  echo "== CS=$cs"
  PQ_LUT2_CS=$cs timeout 300 python tools/microbench.py --shapes qkv7b,o7b,gateup7b,down7b --formats 8x8x1 --tokens 1 --algos lut2 --pdl 1 --dequant 0 2>&1 | grep -o "'shape': '[a-z0-9]*'.*'us': [0-9.]*"
done
