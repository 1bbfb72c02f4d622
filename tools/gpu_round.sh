set -x
mkdir -p gpurun_out
timeout 900 python tools/microbench.py --shapes gateup7b,down7b,gateup70b --formats 4x4x1,4x8x1,4x16x1,8x4x1,8x8x1,8x16x1,12x4x1,12x8x1,12x16x1 --tokens 1 --algos auto --pdl 1 --dequant 0 > gpurun_out/r1_27_microbench_sweep.jsonl 2>&1; echo "sweep rc=$?"
for cfg in "gateup70b 28672,28672 8192 4" "down7b 4096 11008 8" "gateup7b 11008,11008 4096 8"; do
  set -- $cfg
  A="tools/run_one.py --parts $2 --k $3 --algo auto --nsets $4 --iters 10"
  timeout 120 python $A && timeout 300 ncu --set full --clock-control none --import-source on -k regex:pq_ -s 8 -c 1 -f -o gpurun_out/r1_27_$1 python $A > gpurun_out/r1_27_ncu_$1.log 2>&1
done
ls -la gpurun_out/r1_27*.ncu-rep
python - <<'PY'
import ast
for l in open('gpurun_out/r1_27_microbench_sweep.jsonl'):
    l=l.strip()
    if l.startswith('{'):
        d=ast.literal_eval(l); print(d['shape'],d['fmt'],'t=%d'%d['t'],'us=%.1f gbs=%.0f frac=%.3f'%(d['us'],d['gbs'],d['frac']))
    else: print(l)
PY
