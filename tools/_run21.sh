mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29528 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/r2_scale_f_n8.json 2> gpurun_out/r2_scale_f_n8.err; echo n8 rc=$?
