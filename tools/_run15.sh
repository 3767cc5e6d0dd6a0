python tools/stream_pair_probe.py alone
CUDA_VISIBLE_DEVICES=0 python tools/stream_pair_probe.py pairA & 
CUDA_VISIBLE_DEVICES=1 python tools/stream_pair_probe.py pairB &
wait
CUDA_VISIBLE_DEVICES=0 taskset -c 0-5 python tools/stream_pair_probe.py pinA & 
CUDA_VISIBLE_DEVICES=1 taskset -c 12-17 python tools/stream_pair_probe.py pinB &
wait
nvidia-smi topo -m 2>&1 | head -8
lscpu | grep -i "numa\|socket\|model name" | head
