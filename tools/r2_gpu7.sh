mkdir -p gpurun_out
O=gpurun_out/r2_17
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3 | tee ${O}_pytest_tail.txt
timeout 300 python bench.py --no-cpu --steps 100 > ${O}_bench.json 2>> ${O}.err
python - <<PY
import json
d=json.loads([l for l in open("${O}_bench.json") if l.startswith("{")][-1])
print(round(d["value"],1), round(d["ms_per_step"],4), {k:round(x["us"],2) for k,x in d["kernels"].items()}, d["checked"])
PY
timeout 120 python tools/trace_lut.py --n 12288 --k 4096 --parts 3 --pdl 1 --chain 3 2>&1 | grep -v "max active" | head -12
