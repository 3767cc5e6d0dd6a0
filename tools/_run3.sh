mkdir -p gpurun_out
python bench.py --steps 10 --warmup 3 > gpurun_out/r2_bench1_n1.json 2> gpurun_out/r2_bench1_n1.err; echo n1 rc=$?
tail -3 gpurun_out/r2_bench1_n1.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r2_bench1_n2.json 2> gpurun_out/r2_bench1_n2.err; echo n2 rc=$?
tail -5 gpurun_out/r2_bench1_n2.err
