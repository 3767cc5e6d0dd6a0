"""GPU probe: where the time of the frame-stream + match loop goes (host-side timers)."""
import os, sys, time
sys.path.insert(0, os.environ.get("CLB_ROOT", os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
import coloc_b200 as clb
from coloc_b200 import synth
w, h = 1280, 720
views, _ = synth.config_views(w, h, 4)
pins = []
for i in range(4):
    t = torch.empty((h, w), dtype=torch.uint8).pin_memory(); t.numpy()[:] = views[i]; pins.append(t)
fs = clb.FrameStream(w, h, depth=5, maxkp=120000)
m = clb.Matcher(maxkp=120000)
out = torch.zeros(120000, dtype=torch.int32, device="cuda")
st = torch.cuda.current_stream().cuda_stream
T = {"submit": 0.0, "next": 0.0, "match": 0.0, "cpu": 0.0}
def run(n):
    sub = done = 0; prev = None
    while done < n:
        t0 = time.perf_counter()
        while sub < n and fs.in_flight() < 3:
            fs.submit(pins[sub % 4].numpy(), tag=sub); sub += 1
        t1 = time.perf_counter()
        tag, k, d, dk, dd, cnt = fs.next(copy=False)
        t2 = time.perf_counter()
        if prev is not None:
            m.match_dev(dd, cnt, prev[0], prev[1], 40, out, stream=st)
            t3 = time.perf_counter()
            out[:cnt].cpu()
            t4 = time.perf_counter()
            T["match"] += t3 - t2; T["cpu"] += t4 - t3
        T["submit"] += t1 - t0; T["next"] += t2 - t1
        prev = (dd, cnt); done += 1
run(8)
for k in T: T[k] = 0.0
torch.cuda.synchronize(); t0 = time.perf_counter(); run(256); torch.cuda.synchronize(); dt = time.perf_counter() - t0
print(f"{256 / dt:.0f} frames/s; per frame us:", {k: round(v / 256 * 1e6, 1) for k, v in T.items()})
clb.set_timing(True)
run(8)
print("kernels ms:", {k: round(clb.last_kernel_ms(k), 4) for k in ("match", "merge", "expand", "describe", "fast")})
