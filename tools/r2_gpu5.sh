mkdir -p gpurun_out
O=gpurun_out/r2_15
for shp in "--n 12288 --k 4096 --parts 3" "--n 4096 --k 11008 --parts 1" "--n 22016 --k 4096 --parts 2"; do
  for pc in "--pdl 0 --chain 0" "--pdl 1 --chain 3"; do
    echo "== $shp $pc"
    timeout 120 python tools/trace_lut.py $shp $pc 2>&1 | grep -v "max active"
  done
done > ${O}_traces.txt 2>&1
cat ${O}_traces.txt
