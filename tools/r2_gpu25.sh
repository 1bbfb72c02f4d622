mkdir -p gpurun_out
O=gpurun_out/r2_32
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3 | tee ${O}_pytest_tail.txt
for i in 1 2; do
timeout 300 python bench.py --no-cpu --no-extras --steps 100 > ${O}_bench.json 2>> ${O}.err
python - <<PY
import json
d=json.loads([l for l in open("${O}_bench.json") if l.startswith("{")][-1])
print(round(d["value"],1), round(d["ms_per_step"],4), {k:round(x["us"],2) for k,x in d["kernels"].items()}, d["checked"])
PY
done
