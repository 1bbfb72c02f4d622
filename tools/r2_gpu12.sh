mkdir -p gpurun_out
O=gpurun_out/r2_22
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 | tee ${O}_pytest.txt
timeout 900 python tools/microbench.py --shapes o7b,down7b,qkv7b,down70b --formats 8x8x1,8x8x2,16x8x1 --tokens 16,32,64 --dequant 0 --algos tc,mma --ref 1 --warm 0.1 2>&1 | grep "'op'" | python -c "
import sys,ast
for l in sys.stdin:
    r=ast.literal_eval(l); print(r['shape'],r['fmt'],r['t'],r['algo'],round(r['us'],1),'us')" | tee ${O}_t16_64.txt
timeout 600 python tools/microbench.py --shapes gateup8b,down8b --formats 8x8x2 --tokens 2048 --dequant 0 --algos tc 2>&1 | grep "'op'" | cut -c1-200
