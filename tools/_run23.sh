mkdir -p gpurun_out
CLB_LIB=lab python tools/tc_trace.py 20480 2>&1 | grep -v "^tc4" > gpurun_out/r2_tc_trace.log; tail -6 gpurun_out/r2_tc_trace.log
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fuzz.py -m gpu -q -x -k "match or tc or pairs or shapes" 2>&1 | tail -2
CLB_LIB=lab timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "match or tc or pairs" 2>&1 | tail -2
