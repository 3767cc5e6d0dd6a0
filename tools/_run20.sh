mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r2_try2_n2.json 2> gpurun_out/r2_try2_n2.err; echo n2 rc=$?
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2_try2_n2.json').read().strip().splitlines()[-1])
for k,v in d['extra'].items():
    if 'frames_per_s' in v: print(k, round(v['frames_per_s']), v['wall_ms'], v['frames_done_ms'])
PY
