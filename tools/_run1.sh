set -x
mkdir -p gpurun_out
nvidia-smi -L
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_smoke.log 2>&1; echo smoke rc=$?
tail -3 gpurun_out/r2_smoke.log
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_tests.log 2>&1; echo tests rc=$?
tail -15 gpurun_out/r2_tests.log
CLB_LIB=lab timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "fast or match_vs or tc_vs or knobs or describe" > gpurun_out/r2_tests_lab.log 2>&1; echo labtests rc=$?
tail -5 gpurun_out/r2_tests_lab.log
python tools/stage_probe.py > gpurun_out/r2_stage_probe0.log 2>&1; cat gpurun_out/r2_stage_probe0.log
python bench.py --steps 5 --warmup 3 > gpurun_out/r2_bench0.json 2> gpurun_out/r2_bench0.err; echo bench rc=$?
head -c 1500 gpurun_out/r2_bench0.json
