set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_lut_tma_kgroups.py -x -q -k "lut_tma" 2>&1 | tail -3
timeout 600 python tools/microbench.py --shapes qkv7b,o7b,gateup7b,down7b,gateup70b,down70b,down7b_tp4,down7b_tp8 --formats 8x8x1 --tokens 1 --algos auto --pdl 1 --dequant 0 > gpurun_out/r1_29_microbench.jsonl 2>&1; echo "microbench rc=$?"
python - <<'PY'
import ast
for l in open('gpurun_out/r1_29_microbench.jsonl'):
    l=l.strip()
    if l.startswith('{'):
        d=ast.literal_eval(l); print(d['shape'],d['fmt'],d['algo'],'us=%.1f gbs=%.0f'%(d['us'],d['gbs']))
    else: print(l)
PY
