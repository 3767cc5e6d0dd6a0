mkdir -p gpurun_out
TAG=${1:-d}
for N in 8 4 2; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2952$N bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/r2_scale_${TAG}_n$N.json 2> gpurun_out/r2_scale_${TAG}_n$N.err; echo n$N rc=$?
done
python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r2_scale_${TAG}_n1.json 2> gpurun_out/r2_scale_${TAG}_n1.err; echo n1 rc=$?
