mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r2_try_n2.json 2> gpurun_out/r2_try_n2.err; echo rc=$?
CUDA_DEVICE_MAX_CONNECTIONS=32 python tools/stream_pair_probe.py conn32
