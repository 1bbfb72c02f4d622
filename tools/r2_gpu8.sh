mkdir -p gpurun_out
O=gpurun_out/r2_18
timeout 900 python -m pytest tests/test_gpu_fused.py -x -q 2>&1 | tail -12 | tee ${O}_pytest_fused.txt
timeout 300 python tools/microbench.py --shapes down7b_tp4,down7b_tp8 --formats 8x8x1 --tokens 1 --dequant 0 --algos auto,lut2 2>&1 | tail -5
