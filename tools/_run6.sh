mkdir -p gpurun_out
for t in detect4k detect; do
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r2_launches_$t.csv python tools/prof_target.py $t > gpurun_out/r2_ncu_$t.log 2>&1
python tools/summarize_launches.py gpurun_out/r2_launches_$t.csv gpurun_out/r2_launches_$t.md "$t" | tail -12
done
ncu --set full --clock-control none --import-source on -k regex:"resize_levels|fast_nms2|compact2" -c 6 -o gpurun_out/r2_detect4k_a python tools/prof_target.py detect4k > gpurun_out/r2_ncu_full_a.log 2>&1; echo ncu rc=$?
ls -la gpurun_out/*.ncu-rep
