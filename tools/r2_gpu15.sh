mkdir -p gpurun_out
O=gpurun_out/r2_24
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -k "mma" 2>&1 | tail -3
timeout 900 python tools/microbench.py --shapes o7b,down7b,qkv7b,gateup7b,down70b --formats 8x8x2,16x8x1 --tokens 8,16,32,64 --dequant 0 --algos tc,mma --ref 1 --warm 0.1 2>&1 | grep "'op'" | python -c "
import sys,ast
for l in sys.stdin:
    r=ast.literal_eval(l); print(r['shape'],r['fmt'],r['t'],r['algo'],round(r['us'],1),'us')" | tee ${O}_t8_64.txt
