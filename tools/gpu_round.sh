set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -3
timeout 600 python bench.py > gpurun_out/r1_33_bench.json 2> gpurun_out/r1_33_bench.err; echo "bench rc=$?"; cut -c1-420 gpurun_out/r1_33_bench.json
