set -x
mkdir -p gpurun_out
N=$1
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tools/tp_check.py > gpurun_out/r1_23_tp_check_$N.log 2>&1; echo "tp_check rc=$?"
grep -E "tp$N|TP_CHECK|Error|error" gpurun_out/r1_23_tp_check_$N.log | tail -8
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 30 --warmup 5 > gpurun_out/r1_23_bench_tp$N.json 2> gpurun_out/r1_23_bench_tp$N.err; echo "bench tp rc=$?"
python -c "
import json,sys
d=json.loads([l for l in open('gpurun_out/r1_23_bench_tp$N.json') if l.startswith('{')][0]); print(d['n_gpus'], d['value'], d['ms_per_step'], d['e2e']['value']); print({k:(round(v['us'],2),round(v['gbs'])) for k,v in d['kernels'].items()})"
