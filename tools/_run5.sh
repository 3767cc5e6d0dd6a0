mkdir -p gpurun_out
cat > /tmp/t.py <<'PY'
import numpy as np, coloc_b200 as clb
from coloc_b200 import synth
img = synth.textured_frame(640, 480, 5)
det = clb.Detector(640, 480, maxkp=1000)
try:
    got = det.pyramid(img)
    print("ok", [g.shape for g in got])
except Exception as e:
    print("ERR", e)
PY
timeout 300 compute-sanitizer --tool memcheck python /tmp/t.py 2>&1 | head -60
