set -x
mkdir -p gpurun_out
O=gpurun_out/r2_13
N=${1:-2}
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tools/tp_check.py 2>&1 | grep -v "^W\|^\*\*\*" | tail -150 | tee ${O}_tp_check_$N.txt
for fused in 1 0; do
PQ_FUSED_ALLREDUCE=$fused timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 100 > ${O}_bench_tp${N}_fused${fused}.json 2>> ${O}_bench.err
python - <<PY
import json
d=json.loads([l for l in open("${O}_bench_tp${N}_fused${fused}.json") if l.startswith("{")][-1])
print("tp$N fused=$fused", round(d["value"],1), round(d["ms_per_step"],4), {k:round(x["us"],2) for k,x in d["kernels"].items()}, d["checked"], d["check"])
PY
done
tail -5 ${O}_bench.err
