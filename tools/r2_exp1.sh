set -x
mkdir -p gpurun_out
O=gpurun_out/r2_01
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4 | tee ${O}_pytest.txt
for v in "PQ_LUT2_CW=16" "PQ_LUT2_CW=8 PQ_LUT2_PER_SM8=1" "PQ_LUT2_CW=8 PQ_LUT2_PER_SM8=2" "PQ_LUT2_CW=8 PQ_LUT2_PER_SM8=1 PQ_LUT2_CS=8"; do
  tag=$(echo "$v" | tr ' =' '__')
  env $v timeout 300 python bench.py --no-cpu --steps 50 > ${O}_bench_${tag}.json 2> ${O}_bench_${tag}.err; echo "rc=$?"
  python - <<PY
import json
d=json.load(open("${O}_bench_${tag}.json"))
print("$v", d["value"], d["ms_per_step"], {k:round(x["us"],2) for k,x in d["kernels"].items()})
PY
done
for v in "PQ_LUT2_CW=16" "PQ_LUT2_CW=8"; do
  tag=$(echo "$v" | tr ' =' '__')
  for shp in "--n 4096 --k 4096 --parts 1" "--n 4096 --k 11008 --parts 1" "--n 22016 --k 4096 --parts 2" "--n 12288 --k 4096 --parts 3"; do
    for pc in "--pdl 0 --chain 0" "--pdl 1 --chain 3"; do
      echo "== $v $shp $pc"
      env $v timeout 120 python tools/trace_lut.py $shp $pc
    done
  done
done > ${O}_traces.txt 2>&1
tail -5 ${O}_traces.txt
