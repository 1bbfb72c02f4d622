export PQ_B200_LIB=build/exp_tctrace/libpq_b200.so
python tools/trace_tc.py 128 256 32
python tools/trace_tc.py 4096 4096 32
python tools/trace_tc.py 4096 4096 32 8x8x2
python tools/trace_tc.py 4096 4096 64 16x8x1
