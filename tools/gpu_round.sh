set -x
mkdir -p gpurun_out
for dbg in 0 4 8 6; do
  echo "=== DEBUG=$dbg" >> gpurun_out/r1_15_trace_lut2.txt
  PQ_LUT2_DEBUG=$dbg timeout 120 python tools/trace_lut.py --n 22016 --k 4096 --parts 2 --algo 5 2>&1 | grep -v "max active" >> gpurun_out/r1_15_trace_lut2.txt
done
cat gpurun_out/r1_15_trace_lut2.txt
