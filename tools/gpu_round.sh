set -x
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/tp_check.py > gpurun_out/r1_05_tp_check.log 2>&1; echo "tp_check rc=$?"
grep -E "tp2|TP_CHECK|rank" gpurun_out/r1_05_tp_check.log | tail -8
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 30 --warmup 5 > gpurun_out/r1_05_bench_tp2.json 2> gpurun_out/r1_05_bench_tp2.err; echo "bench tp2 rc=$?"
PQ_ONESHOT_ALLREDUCE=0 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --steps 30 --warmup 5 > gpurun_out/r1_05_bench_tp2_nccl.json 2> gpurun_out/r1_05_bench_tp2_nccl.err; echo "bench tp2 nccl rc=$?"
