mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "pyramid or fast or describe or golden or frame or config or pair_pipeline or pinned" > gpurun_out/r2_tests_b.log 2>&1; echo tests rc=$?
tail -4 gpurun_out/r2_tests_b.log
CLB_LIB=lab timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "fast" > gpurun_out/r2_tests_lab_b.log 2>&1; echo labtests rc=$?
tail -3 gpurun_out/r2_tests_lab_b.log
python tools/stage_probe.py 2>&1 | tee gpurun_out/r2_stage_probe3.log
for t in detect4k; do
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r2_launches_$t.csv python tools/prof_target.py $t > gpurun_out/r2_ncu_$t.log 2>&1
python tools/summarize_launches.py gpurun_out/r2_launches_$t.csv gpurun_out/r2_launches_$t.md "$t" | tail -9
done
