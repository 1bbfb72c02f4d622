mkdir -p gpurun_out
O=gpurun_out/r2_25
timeout 900 python tools/microbench.py --shapes o7b,down7b,qkv7b,gateup7b,gateup70b,down70b --formats 16x8x1 --tokens 1,2,4,8 --dequant 0 --algos tc,direct --ref 1 --warm 0.1 2>&1 | grep "'op'" | python -c "
import sys,ast
for l in sys.stdin:
    r=ast.literal_eval(l); print(r['shape'],r['fmt'],r['t'],r['algo'],round(r['us'],1),'us')" | tee ${O}_1x16_small_t.txt
timeout 900 python tools/microbench.py --shapes o7b,down7b,qkv7b,gateup7b,down70b --formats 8x8x1 --tokens 8,16,32,64 --dequant 0 --algos tc,mma --warm 0.1 2>&1 | grep "'op'" | python -c "
import sys,ast
for l in sys.stdin:
    r=ast.literal_eval(l); print(r['shape'],r['fmt'],r['t'],r['algo'],round(r['us'],1),'us')" | tee ${O}_8x8x1_t8_64.txt
