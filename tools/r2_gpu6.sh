mkdir -p gpurun_out
O=gpurun_out/r2_16
for v in "" build/exp_NO_HMMA/libpq_b200.so build/exp_NO_STS/libpq_b200.so; do
  PQ_B200_LIB=$v timeout 300 python bench.py --no-cpu --steps 100 > ${O}_tmp.json 2>> ${O}.err
  python - <<PY
import json
d=json.loads([l for l in open("${O}_tmp.json") if l.startswith("{")][-1])
print("variant '$v'", round(d["value"],1), round(d["ms_per_step"],4), {k:round(x["us"],2) for k,x in d["kernels"].items()}, d["checked"])
PY
done | tee ${O}_build_cost.txt
