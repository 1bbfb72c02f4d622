mkdir -p gpurun_out
O=gpurun_out/r2_05
for v in "PQ_LUT2_CW=16" "PQ_LUT2_CW=8" "PQ_LUT2_CW=8 PQ_LUT2_EARLY_LD=1"; do
  for mask in 0x185 0xffff; do
  for shp in "--n 4096 --k 4096 --parts 1" "--n 22016 --k 4096 --parts 2"; do
    for pdl in 1 0; do
      echo "== $v $shp pdl=$pdl mask=$mask"
      env $v PQ_LUT2_TRACE_MASK=$mask timeout 120 python tools/trace_chain.py $shp --pdl $pdl
    done
  done
  done
done > ${O}_chain.txt 2>&1
