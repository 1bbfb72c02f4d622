set -x
mkdir -p gpurun_out
O=gpurun_out/r2_27
N=${1:-2}
BATCHES=${2:-"1 2048"}
[ "$3" = "notp" ] || timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tools/tp_check.py 2>&1 | grep -v "^W\|^\*\*\*\| ok$" | tail -12 | tee ${O}_tp_check_$N.txt
for B in $BATCHES; do
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 20 --model llama-2-70b --batch $B > ${O}_bench70b_tp${N}_b${B}.json 2>> ${O}_bench.err
python - <<PY
import json
try:
    d=json.loads([l for l in open("${O}_bench70b_tp${N}_b${B}.json") if l.startswith("{")][-1])
    print("70b tp$N batch $B", round(d["value"],1), "tok/s", round(d["ms_per_step"],4), "ms", {k:round(x["us"],2) for k,x in d["kernels"].items()}, d["checked"], d["config"]["allreduce"], d["config"]["launch"])
except Exception as e:
    print("70b tp$N batch $B failed", e)
PY
done
[ "$4" = "7b" ] && { timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $N --steps 100 > ${O}_bench7b_tp${N}.json 2>> ${O}_bench.err
python - <<PY
import json
d=json.loads([l for l in open("${O}_bench7b_tp${N}.json") if l.startswith("{")][-1])
print("7b tp$N", round(d["value"],1), "tok/s", round(d["ms_per_step"],4), "ms", {k:round(x["us"],2) for k,x in d["kernels"].items()}, d["checked"], d["config"]["allreduce"])
PY
}
tail -5 ${O}_bench.err | cut -c1-300
