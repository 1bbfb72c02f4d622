set -x
mkdir -p gpurun_out
O=gpurun_out/r2_11
timeout 600 python -m pytest tests/test_gpu_fused.py -x -q 2>&1 | tail -15 | tee ${O}_pytest_fused.txt
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4 | tee ${O}_pytest_tail.txt
timeout 300 python bench.py --no-cpu > ${O}_bench.json 2> ${O}_bench.err; echo "bench rc=$?"; cut -c1-300 ${O}_bench.json
python - <<PY
import json
d=json.load(open("${O}_bench.json"))
print(d["value"], d["ms_per_step"], {k:round(x["us"],2) for k,x in d["kernels"].items()}, d["checked"])
PY
