set -x
mkdir -p gpurun_out
run() { timeout 600 python tools/microbench.py --shapes gateup7b,down7b,gateup70b --formats 8x8x1 --tokens 1 --algos auto --pdl 1 --dequant 0 2>&1 | grep -o "'shape': '[a-z0-9]*'\|'us': [0-9.]*" | paste - - ; }
echo "== new builder"; run
cp vllm_pq_b200/libpq_b200.so /tmp/new.so; cp gpurun_out_old_libpq_b200.so vllm_pq_b200/libpq_b200.so
echo "== old builder"; run
cp /tmp/new.so vllm_pq_b200/libpq_b200.so
echo "== new builder again"; run
