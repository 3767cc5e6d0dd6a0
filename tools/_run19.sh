mkdir -p gpurun_out
./tools/probe/lds_probe > gpurun_out/lds_probe_plain.log 2>&1; echo plain rc=$?; tail -1 gpurun_out/lds_probe_plain.log
ncu --metrics l1tex__data_pipe_lsu_wavefronts_mem_shared_op_ld.sum,smsp__inst_executed_op_shared_ld.sum,l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_ld.sum --clock-control none -c 500 --csv --log-file gpurun_out/lds_probe_ncu.csv ./tools/probe/lds_probe > gpurun_out/lds_probe_ncu.log 2>&1; echo ncu rc=$?
tail -2 gpurun_out/lds_probe_ncu.log
