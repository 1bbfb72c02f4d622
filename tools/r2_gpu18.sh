mkdir -p gpurun_out
O=gpurun_out/r2_26
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 | tee ${O}_pytest.txt


import json
d=json.loads([l for l in open("${O}_bench.json") if l.startswith("{")][-1])
print(round(d["value"],1), round(d["ms_per_step"],4), {k:round(x["us"],2) for k,x in d["kernels"].items()}, d["checked"])
print(d["batch"]); print({k:round(v["tflops"]) for k,v in d["prefill"].items() if isinstance(v,dict) and "tflops" in v})
PY
