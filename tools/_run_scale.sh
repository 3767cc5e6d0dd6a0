mkdir -p gpurun_out
TAG=${1:-a}
for N in 8 4 2; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$N bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/r2_scale_${TAG}_n$N.json 2> gpurun_out/r2_scale_${TAG}_n$N.err; echo n$N rc=$?
done
python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r2_scale_${TAG}_n1.json 2> gpurun_out/r2_scale_${TAG}_n1.err; echo n1 rc=$?
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533 tools/e2e_timeline.py > gpurun_out/r2_timeline_n8.json 2> gpurun_out/r2_timeline_n8.err; echo tl8 rc=$?
python tools/e2e_timeline.py > gpurun_out/r2_timeline_n1.json 2>/dev/null; echo tl1 rc=$?
timeout 600 python -m pytest tests/test_sharded_gpu.py -m gpu -q 2>&1 | tail -3
