#!/usr/bin/env python3
"""GPU exploration run (not a test, not the bench): pipe micro-benchmarks, quick parity against
the oracle, and timing sweeps over the match kernel's knobs.  Writes gpurun_out/explore.json."""
import json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import coloc_b200 as clb
from coloc_b200 import synth
import oracle as O

out = {}
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)


def dump():
    with open(os.path.join(ROOT, "gpurun_out", "explore.json"), "w") as f:
        json.dump(out, f, indent=1)


print(torch.cuda.get_device_name(0))
out["pipes"] = clb.microbench_pipes(0)
print("pipes", out["pipes"]); dump()

# ---- match parity + sweep
m = clb.Matcher(maxkp=1 << 16)
rng = np.random.default_rng(1)
par = {}
for (nq, nt) in [(1, 1), (5, 3), (300, 1000), (1000, 777), (2500, 4100)]:
    T = synth.random_descriptors(nt, 10 + nt); Q, _ = synth.planted_queries(T, nq, 11 + nq)
    Q[: nq // 4] = T[rng.integers(0, nt, nq // 4)]   # exact duplicates -> ties
    for hs in range(4):
        clb.set_param("match_hs", hs)
        got = m.compute_matches(Q, T, 40, want_dist=True)
        ref = O.match_2nn(Q, T, 40)
        par[f"{nq}x{nt}_hs{hs}"] = bool(all(np.array_equal(a, b) for a, b in zip(got, ref)))
out["match_parity"] = par
print("match parity", par); dump()

clb.set_timing(True)
sweep = {}
nq, nt = 50000, 200000
T = torch.from_numpy(synth.random_descriptors(nt, 42).view(np.int64)).cuda()
Q = torch.from_numpy(synth.random_descriptors(nq, 43).view(np.int64)).cuda()
dm = torch.zeros(nq, dtype=torch.int32, device="cuda")
for rq in (1, 2, 4):
    for hs in (0, 1, 2, 3):
        clb.set_param("match_rq", rq); clb.set_param("match_hs", hs)
        ts = []
        for it in range(4):
            m.match_dev(Q, nq, T, nt, 40, dm)
            torch.cuda.synchronize()
            ts.append(clb.last_kernel_ms("match"))
        best = min(ts[1:])
        sweep[f"rq{rq}_hs{hs}"] = {"ms": best, "gcmp_s": nq * nt / best / 1e6}
        print(rq, hs, sweep[f"rq{rq}_hs{hs}"])
out["match_sweep_50k_x_200k"] = sweep; dump()
clb.set_param("match_rq", 0); clb.set_param("match_hs", 2)

# ---- detector parity + timing (C2-like frame)
for (w, h) in [(640, 480), (1280, 720)]:
    a, b, H = synth.frame_pair(w, h, 0)
    det = clb.Detector(w, h, maxkp=60000)
    okps, odesc, opyr = O.detect_describe(a, want_pyramid=True)
    pyr = det.pyramid(a)
    pd = [int(np.abs(x.astype(int) - y.astype(int)).max()) for x, y in zip(pyr, opyr)]
    kps, desc = det.detect_and_describe(a)
    res = {"pyr_maxdiff": pd, "n": int(len(kps)), "n_oracle": int(len(okps)),
           "kps_equal": bool(len(kps) == len(okps) and np.array_equal(kps, okps)),
           "desc_equal": bool(desc.shape == odesc.shape and np.array_equal(desc, odesc))}
    if not res["kps_equal"] and len(kps) == len(okps):
        for f in ("x", "y", "score", "angle", "scale"):
            res["mismatch_" + f] = int((kps[f] != okps[f]).sum())
    # stage isolation: describe with the oracle's pyramid + keypoints
    d2 = det.describe(opyr, okps)
    res["describe_stage_equal"] = bool(np.array_equal(d2, odesc))
    if not res["describe_stage_equal"]:
        res["describe_bad_rows"] = int((d2 != odesc).any(axis=1).sum())
    ts = {}
    for it in range(5):
        t0 = time.perf_counter(); det.detect_and_describe(a); t1 = time.perf_counter()
        ts = {k: clb.last_kernel_ms(k) for k in ("resize", "fast", "describe")}
        ts["host_ms"] = (t1 - t0) * 1e3
    res["timing"] = ts
    out[f"detect_{w}x{h}"] = res
    print(w, h, res); dump()
    det.close()
print("launches", clb.launch_count())
