mkdir -p gpurun_out
O=gpurun_out/r2_23
for ks in 1 4 8; do
echo "== forced ksplit <= $ks"
PQ_TC_KSPLIT=$ks timeout 300 python tools/microbench.py --shapes o7b,down7b --formats 8x8x1,8x8x2 --tokens 32 --dequant 0 --algos tc --warm 0.1 2>&1 | grep "'op'" | python -c "
import sys,ast
for l in sys.stdin:
    r=ast.literal_eval(l); print(r['shape'],r['fmt'],r['t'],r['algo'],round(r['us'],1),'us')"
done | tee ${O}_ksplit.txt
