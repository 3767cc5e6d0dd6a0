mkdir -p gpurun_out
python tools/stage_probe.py 2>&1 | tee gpurun_out/r2_stage_probe_b.log
timeout 1500 python -m pytest tests -m gpu -q -x 2>&1 | tail -4 | tee gpurun_out/r2_tests_b.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
python bench.py > gpurun_out/r2_bench_default.json 2> gpurun_out/r2_bench_default.err; echo bench rc=$?
