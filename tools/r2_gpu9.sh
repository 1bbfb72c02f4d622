mkdir -p gpurun_out
O=gpurun_out/r2_19
( time timeout 600 python bench.py > ${O}_bench.json 2> ${O}_bench.err ) 2>&1 | tail -3
tail -5 ${O}_bench.err
python - <<PY
import json
d=json.loads([l for l in open("${O}_bench.json") if l.startswith("{")][-1])
print(round(d["value"],1), round(d["ms_per_step"],4), {k:round(x["us"],2) for k,x in d["kernels"].items()}, d["checked"])
for k in ("batch","gemv_single","sweep","llama70b_layer","prefill","ref_cuda"):
    print(k, json.dumps(d.get(k))[:1500])
PY
( time timeout 600 python bench.py --impl reference --steps 20 --warmup 2 ) 2>&1 | cut -c1-300 | tail -5
