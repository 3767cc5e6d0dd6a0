mkdir -p gpurun_out
N=$1
if [ "$N" = "1" ]; then
python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r2_scale_g_n1.json 2> gpurun_out/r2_scale_g_n1.err; echo n1 rc=$?
else
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2953$N bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/r2_scale_g_n$N.json 2> gpurun_out/r2_scale_g_n$N.err; echo n$N rc=$?
fi
