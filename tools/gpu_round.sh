# One round's GPU validation on a single B200 (gpurun --timeout 1800 -- 'bash tools/gpu_round.sh rN_MM'):
# full GPU suite, smoke, bench (+ reference arm), the library's launch list under ncu, one --set full
# capture of the dominant decode kernel and of the prefill GEMM.  Outputs under gpurun_out/<tag>_*.
TAG=${1:-round}
mkdir -p gpurun_out
O=gpurun_out/${TAG}
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3 | tee ${O}_pytest_gpu_tail.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -6 | tee ${O}_smoke.txt
timeout 600 python bench.py > ${O}_bench.json 2> ${O}_bench.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 20 --warmup 2 > ${O}_bench_reference.json 2>> ${O}_bench.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:pq_ -c 600 --csv \
  --log-file ${O}_ncu_launches_bench.csv python bench.py --steps 2 --warmup 1 --no-cpu --no-extras > ${O}_ncu_bench.log 2>&1
for spec in "down7b 4096 11008 8x8x1 1 lut2" "gateup70b 28672,28672 8192 8x8x1 1 lut2" "gateup8b 14336,14336 4096 8x8x1 2048 tc"; do
  set -- $spec
  timeout 600 ncu --set full --import-source on --clock-control none -k regex:pq_ge -s 2 -c 1 -f -o ${O}_ncu_$1_$4_t$5 \
    python tools/run_one.py --parts $2 --k $3 --fmt $4 --t $5 --algo $6 --iters 4 > /dev/null 2>&1
done
