#!/usr/bin/env python3
"""The reference's own CUDA kernels (oracle/_ref/libcoloc_ref_gpu.so: CLATCH.cu, CUDAK2NN.cu compiled unmodified for
sm_100a) timed on the B200 next to the new kernels, same inputs, outputs compared bit for bit.  This is the
"kernel to beat on the same hardware" column of BASELINE.md section 4 -- test infrastructure, not part of bench.py.
usage (GPU box): python tools/ref_gpu_timing.py > gpurun_out/ref_gpu_timing.json"""
import json, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import coloc_b200 as clb
import oracle as O
from coloc_b200 import synth

out = {}
clb.set_timing(True)
# ---- describe: 1280x720 config frame (~20k keypoints) and 3840x2160 (~50k)
for (w, h) in ((1280, 720), (3840, 2160)):
    img = synth.config_frame(w, h, 0)
    det = clb.Detector(w, h, maxkp=400000)
    for _ in range(3):
        kps, desc = det.detect_and_describe(img)
    ours_ms = clb.last_kernel_ms("describe")
    pyr = det.pyramid(img)
    ref, ref_ms = O.ref_gpu_latch(pyr, kps, iters=5)
    out[f"describe_{w}x{h}"] = {"keypoints": int(len(kps)), "reference_CLATCH_ms": ref_ms, "latch_kernel_ms": ours_ms,
                                "speedup": ref_ms / ours_ms, "bit_exact": bool(np.array_equal(ref, desc))}
    det.close()
# ---- match: pair-sized and map-sized
m = clb.Matcher(maxkp=1000)
for (nq, nt) in ((20000, 20000), (50000, 200000)):
    T = synth.random_descriptors(nt, 42); Q, _ = synth.planted_queries(T, nq, 43)
    ref, ref_ms = O.ref_gpu_k2nn(Q, T, 40, iters=3)
    res = {}
    for name, tc in (("popc", 0), ("tensor", 1)):
        clb.set_param("match_tc", tc)
        for _ in range(3):
            got = m.compute_matches(Q, T, 40)
        res[name] = (clb.last_kernel_ms("match") + clb.last_kernel_ms("merge"), bool(np.array_equal(got, ref)))
    clb.set_param("match_tc", -1)
    out[f"match_{nq}x{nt}"] = {"reference_CUDAK2NN_ms": ref_ms, "reference_cmp_per_s": nq * nt / ref_ms * 1e3,
                               "match_kernel_popc_ms": res["popc"][0], "match_tc2_kernel_ms": res["tensor"][0],
                               "tensor_cmp_per_s": nq * nt / res["tensor"][0] * 1e3,
                               "speedup_popc": ref_ms / res["popc"][0], "speedup_tensor": ref_ms / res["tensor"][0],
                               "bit_exact": res["popc"][1] and res["tensor"][1]}
print(json.dumps(out, indent=1))
