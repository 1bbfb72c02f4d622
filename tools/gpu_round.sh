set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/r1_20_pytest_gpu.log 2>&1; echo "pytest rc=$?"
tail -4 gpurun_out/r1_20_pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -4
timeout 600 python bench.py > gpurun_out/r1_20_bench.json 2> gpurun_out/r1_20_bench.err; echo "bench rc=$?"
cat gpurun_out/r1_20_bench.json
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu > /dev/null 2>&1 && timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:pq_ -c 1200 --csv --log-file gpurun_out/r1_20_ncu_launches_bench.csv python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/r1_20_ncu_launches.log 2>&1; echo "ncu launches rc=$?"
tail -2 gpurun_out/r1_20_ncu_launches.log
