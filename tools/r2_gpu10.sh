mkdir -p gpurun_out
O=gpurun_out/r2_20
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_configs.py -x -q -k "tc or prefill or config3 or golden" 2>&1 | tail -5 | tee ${O}_pytest_tc.txt
for g in 1 2; do
echo "== groups $g"
PQ_TC_GROUPS=$g timeout 600 python tools/microbench.py --shapes gateup8b,down8b --formats 8x8x1,8x8x2,16x8x1 --tokens 2048 --dequant 0 --algos tc 2>&1 | grep "'op'" | python -c "
import sys,ast
for l in sys.stdin:
    r=ast.literal_eval(l); print(r['shape'],r['fmt'],r['t'],round(r['us'],1),'us',round(r['tflops'],1),'TF',r['sm_mhz'])"
done | tee ${O}_prefill_groups.txt
PQ_TC_GROUPS=2 timeout 600 python tools/microbench.py --shapes o7b,down7b --formats 8x8x1,8x8x2 --tokens 32,64,256 --dequant 0 --algos tc 2>&1 | grep "'op'" | python -c "
import sys,ast
for l in sys.stdin:
    r=ast.literal_eval(l); print(r['shape'],r['fmt'],r['t'],round(r['us'],1),'us',round(r['tflops'],1),'TF')"
