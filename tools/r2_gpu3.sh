set -x
mkdir -p gpurun_out
O=gpurun_out/r2_12
timeout 600 python -m pytest tests/test_gpu_fused.py -x -q 2>&1 | tail -15 | tee ${O}_pytest_fused.txt
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4 | tee ${O}_pytest_tail.txt
for i in 1; do
  for v in old new; do
    if [ $v = old ]; then B=build/old_tree/bench.py; else B=bench.py; fi
    timeout 300 python $B --no-cpu --steps 100 > ${O}_bench_${v}_$i.json 2>> ${O}_bench.err
    python - <<PY
import json
d=json.load(open("${O}_bench_${v}_$i.json"))
print("$v $i", round(d["value"],1), round(d["ms_per_step"],4), {k:round(x["us"],2) for k,x in d["kernels"].items()}, d["checked"])
PY
  done
done 2>&1 | grep -v "^+" | tee ${O}_ab.txt
