mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/r2_topo8.txt 2>&1
lscpu | grep -i "numa\|socket\|model name\|^CPU(s)\|thread" >> gpurun_out/r2_topo8.txt
python - >> gpurun_out/r2_topo8.txt 2>&1 <<'PY'
import os, pynvml
pynvml.nvmlInit()
n = pynvml.nvmlDeviceGetCount()
words = ((os.cpu_count() or 64) + 63) // 64
for i in range(n):
    h = pynvml.nvmlDeviceGetHandleByIndex(i)
    try:
        m = pynvml.nvmlDeviceGetCpuAffinity(h, words)
        print("gpu", i, "cpu affinity words", [hex(int(x)) for x in m])
    except Exception as e:
        print("gpu", i, "affinity error", e)
print("allowed", sorted(os.sched_getaffinity(0)))
PY
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/r2_scale_d_n8.json 2> gpurun_out/r2_scale_d_n8.err; echo n8 rc=$?
timeout 600 python -m pytest tests/test_sharded_gpu.py -m gpu -q 2>&1 | tail -3
