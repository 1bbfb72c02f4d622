PQ_TC_KSPLIT=1 timeout 300 python tools/microbench.py --shapes n4096k1024,n4096k256,n512k4096,n128k256,o7b --formats 8x8x1 --tokens 32 --dequant 0 --algos tc,mma --warm 0.1 2>&1 | grep "'op'" | python -c "
import sys,ast
for l in sys.stdin:
    r=ast.literal_eval(l); print(r['shape'],r['fmt'],r['t'],r['algo'],round(r['us'],1),'us')"
