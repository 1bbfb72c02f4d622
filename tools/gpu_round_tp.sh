# Multi-GPU validation (gpurun --gpus N --timeout 1200 -- 'bash tools/gpu_round_tp.sh N rN_MM "1 64 2048"'):
# tools/tp_check.py (every TP linear against the oracle, all-reduce kernels against NCCL, graph replay),
# the Llama-2-7B decode bench at TP = N and Llama-2-70B at the given batch sizes.
N=${1:-2}; TAG=${2:-round}; BATCHES=${3:-"1 2048"}
mkdir -p gpurun_out
O=gpurun_out/${TAG}
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 900 $RUN --master-port 29511 tools/tp_check.py 2>&1 | grep -v "^W\|^\*\*\*" | tail -150 | tee ${O}_tp_check_$N.txt
timeout 600 $RUN --master-port 29512 bench.py --gpus $N --steps 100 2>> ${O}_bench.err | grep "^{" > ${O}_bench7b_tp$N.json
for B in $BATCHES; do
  timeout 900 $RUN --master-port 29513 bench.py --gpus $N --steps 20 --model llama-2-70b --batch $B 2>> ${O}_bench.err | grep "^{" > ${O}_bench70b_tp${N}_b$B.json
done
for f in ${O}_bench7b_tp$N.json ${O}_bench70b_tp${N}_b*.json; do python - <<PY
import json
d = json.load(open("$f"))
print("$f", round(d["value"], 1), d["unit"], round(d["ms_per_step"], 4), "ms", d["checked"], d["config"]["allreduce"])
PY
done
