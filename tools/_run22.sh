mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fuzz.py tests/test_sharded_gpu.py -m gpu -q -x -k "match or tc or pairs or filtered or partial or sharded or shapes or streamed or resident" 2>&1 | tail -3
CLB_LIB=lab python tools/tc_phases.py 2>&1 | grep -v "^tc4"
CLB_LIB=lab python tools/tc_trace.py 20480 2>&1 | grep -v "^tc4" | tail -6
python tools/ref_gpu_timing.py 2>&1 | tail -4
