mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q 2>&1 | tail -5; echo alone rc=$?
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_tests.log 2>&1; echo all rc=$?
tail -7 gpurun_out/r2_tests.log
