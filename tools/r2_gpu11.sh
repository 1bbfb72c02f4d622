mkdir -p gpurun_out
O=gpurun_out/r2_21
for fmt in 8x8x2 16x8x1; do
timeout 300 python tools/run_one.py --parts 14336,14336 --k 4096 --fmt $fmt --t 2048 --algo tc --iters 4 > /dev/null 2>&1 && \
timeout 600 ncu --set full --import-source on --clock-control none -k regex:pq_gemm_tc -s 2 -c 1 -o ${O}_ncu_tc_gateup8b_${fmt} -f python tools/run_one.py --parts 14336,14336 --k 4096 --fmt $fmt --t 2048 --algo tc --iters 4 2>&1 | tail -2
done
ls -la gpurun_out/
