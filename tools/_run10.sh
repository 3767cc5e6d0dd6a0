mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r2_tests.log 2>&1; echo tests rc=$?
tail -6 gpurun_out/r2_tests.log
CLB_LIB=lab timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "fast or match_vs or tc_vs or knobs or describe" > gpurun_out/r2_tests_lab.log 2>&1; echo labtests rc=$?
tail -3 gpurun_out/r2_tests_lab.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
