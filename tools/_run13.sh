mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_sharded_gpu.py tests/test_gpu_parity.py -m gpu -q -x -k "sharded or streamed or match_map or resident or partial" 2>&1 | tail -3
python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r2_bench_n1.json 2> gpurun_out/r2_bench_n1.err; echo bench rc=$?
python bench.py --impl reference --gpus 1 --steps 3 --warmup 1 > gpurun_out/r2_bench_reference_arm.json 2>/dev/null; echo ref rc=$?
ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/r2_launches_bench.csv python bench.py --steps 2 --warmup 1 > gpurun_out/r2_ncu_bench.log 2>&1; echo ncu-bench rc=$?
python tools/summarize_launches.py gpurun_out/r2_launches_bench.csv gpurun_out/r2_launches_bench.md "bench.py --steps 2 --warmup 1 (first 300 launches)" | tail -14
ncu --set full --clock-control none --import-source on -k regex:match_tc4_kernel -s 2 -c 1 -o gpurun_out/r2_match_full python tools/prof_target.py matchfull > gpurun_out/r2_ncu_match_full.log 2>&1; echo ncu-match rc=$?
ncu --set full --clock-control none --import-source on -k regex:"resize_levels|fast_nms2|compact2|fast_xend|latch_kernel" -s 5 -c 5 -o gpurun_out/r2_detect4k python tools/prof_target.py detect4k > gpurun_out/r2_ncu_detect4k_full.log 2>&1; echo ncu-detect rc=$?
for t in detect4k detect; do
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r2_launches_$t.csv python tools/prof_target.py $t > gpurun_out/r2_ncu_$t.log 2>&1
python tools/summarize_launches.py gpurun_out/r2_launches_$t.csv gpurun_out/r2_launches_$t.md "$t" > /dev/null
done
COLOC_B200_NVTX=1 python tools/prof_target.py detect | tail -1
python tools/stage_probe.py 2>&1 | tee gpurun_out/r2_stage_probe.log
ls -la gpurun_out/*.ncu-rep
