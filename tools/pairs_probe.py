"""GPU probe: batched pair list (clb_match_pairs_dev) vs per-pair calls on config-3-like sets; prints times.
python tools/pairs_probe.py"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import coloc_b200 as clb
from coloc_b200 import synth

def main():
    counts = [20744, 14748, 19294, 14453]
    base = synth.random_descriptors(max(counts), 1)
    sets = [synth.planted_queries(base, n, 10 + k)[0] for k, n in enumerate(counts)]
    dev = [torch.from_numpy(s.view(np.int64)).cuda() for s in sets]
    m = clb.Matcher(maxkp=max(counts) + 16); f = clb.FeatureMap()
    st = torch.cuda.current_stream().cuda_stream
    for k in range(4):
        f.put_dev(k, dev[k], counts[k], stream=st)
    pairs = [(i, j) for i in range(4) for j in range(4) if i != j]
    tot = m.pair_rows(f, pairs)
    out_a = torch.zeros(tot, dtype=torch.int32, device="cuda"); out_b = torch.zeros_like(out_a)
    jobs = m._pairs(pairs)
    cmp_step = float(sum(counts[i] * counts[j] for i, j in pairs))

    def calls():
        off = 0
        for i, j in pairs:
            m.match_dev(dev[i], counts[i], dev[j], counts[j], 40, out_a[off:off + counts[i]], stream=st); off += counts[i]

    def lst():
        m.match_pairs_dev(f, jobs, 40, out_b, stream=st)

    for name, fn in (("per-pair calls", calls), ("pair list", lst)):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20):
            fn()
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 20
        print(f"{name}: {ms:.3f} ms/step  {cmp_step / ms / 1e9:.2f} T cmp/s", flush=True)
    print("same:", bool(torch.equal(out_a, out_b)), "matched", int((out_b >= 0).sum()))
    for chunks_tpc in (0,):
        pass

if __name__ == "__main__":
    main()
