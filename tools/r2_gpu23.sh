mkdir -p gpurun_out
O=gpurun_out/r2_30
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:pq_ -c 600 --csv --log-file ${O}_ncu_launches_bench.csv python bench.py --steps 2 --warmup 1 --no-cpu --no-extras > ${O}_ncu_bench.log 2>&1
tail -2 ${O}_ncu_bench.log | cut -c1-200
