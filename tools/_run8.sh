mkdir -p gpurun_out
ncu --set full --clock-control none --import-source on -k regex:"latch_kernel" -s 1 -c 1 -o gpurun_out/r2_latch4k_a python tools/prof_target.py detect4k > gpurun_out/r2_ncu_full_b.log 2>&1; echo ncu rc=$?
