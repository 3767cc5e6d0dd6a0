mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x 2>&1 | tail -4 | tee gpurun_out/r2_tests_c.log
CLB_LIB=lab timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "match or tc or pairs or filtered or partial" 2>&1 | tail -3
CLB_LIB=lab python tools/tc_phases.py 2>&1 | grep -v "^tc4" | tee gpurun_out/r2_tc_phases_b.log
python bench.py --steps 20 --warmup 5 > gpurun_out/r2_bench_e.json 2> gpurun_out/r2_bench_e.err; echo bench rc=$?
