mkdir -p gpurun_out
O=gpurun_out/r2_31
run() { env "$@" timeout 120 python tools/microbench.py --shapes down7b,qkv7b,gateup7b --formats 8x8x1 --tokens 1 --dequant 0 --algos lut2 --pdl 1 --warm 0.1 2>&1 | grep "'op'" | python -c "
import sys,ast
print('$*', ' '.join(f\"{ast.literal_eval(l)['shape']}={ast.literal_eval(l)['us']:.2f}\" for l in sys.stdin))"; }
run A=1
run PQ_LUT2_CW=16 PQ_LUT2_CS=16 PQ_LUT2_KG=1
run PQ_LUT2_CW=16 PQ_LUT2_CS=8 PQ_LUT2_KG=1
run PQ_LUT2_CW=16 PQ_LUT2_CS=8 PQ_LUT2_KG=2
run PQ_LUT2_CW=8 PQ_LUT2_CS=8 PQ_LUT2_KG=3
run PQ_LUT2_CW=8 PQ_LUT2_CS=16 PQ_LUT2_KG=2
run PQ_LUT2_CW=8 PQ_LUT2_PER_SM8=2
run PQ_LUT2_PREFETCH=2
run PQ_LUT2_PREFETCH=16
run PQ_LUT2_STAGES=3
