mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "pyramid or fast or describe or golden or frame or config or pair_pipeline" > gpurun_out/r2_tests_b.log 2>&1; echo tests rc=$?
tail -6 gpurun_out/r2_tests_b.log
for i in 1 2 3; do timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "config5 or config1 or end_to_end" 2>&1 | tail -2; done
python tools/stage_probe.py 2>&1 | tee gpurun_out/r2_stage_probe1.log
