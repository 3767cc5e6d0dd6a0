"""GPU probe: tensor-core match path (match_tc.cu) against the POPC path (match.cu) on the same inputs.
Run on the GPU box:  python tools/tc_probe.py [quick|full]
Prints PASS/FAIL per case and kernel times; exits non-zero on any mismatch."""
import os
import sys
import time
import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import coloc_b200 as clb  # noqa: E402


def rand_rows(rng, n):
    return rng.integers(0, 2**63, size=(n, 8), dtype=np.uint64) ^ (rng.integers(0, 2, size=(n, 8), dtype=np.uint64) << np.uint64(63))


def planted(rng, nQ, nT, flip=0.1):
    T = rand_rows(rng, nT)
    Q = rand_rows(rng, nQ)
    k = int(nQ * 0.8)
    src = rng.integers(0, nT, size=k)
    noise = np.packbits((rng.random((k, 512)) < flip).astype(np.uint8), axis=1, bitorder="little").view(np.uint64)
    Q[:k] = T[src] ^ noise
    return Q, T


def run(m, Q, T, thr, tc):
    clb.set_param("match_tc", tc)
    return m.compute_matches(Q, T, thr, want_dist=True)


def main():
    mode = sys.argv[1] if len(sys.argv) > 1 else "quick"
    global VERS
    VERS = [int(v) for v in sys.argv[2].split(",")] if len(sys.argv) > 2 else [2]
    clb.set_param("match_tc_ver", VERS[-1])
    rng = np.random.default_rng(7)
    m = clb.Matcher(maxkp=1000)
    clb.set_timing(True)
    bad = 0
    cases = [(1, 1), (5, 3), (100, 129), (256, 128), (257, 1000), (300, 5000), (1000, 4096), (5000, 5000), (777, 12345),
             (20000, 20000)]
    if mode == "full":
        cases += [(50000, 200000), (3000, 1000000)]
    for nQ, nT in cases:
        Q, T = planted(rng, nQ, nT)
        if nT >= 8:   # exact duplicates: first-min-wins and second == best on ties
            T[nT // 2] = T[1]; T[nT - 1] = T[1]
            Q[0] = T[1]
        for thr in (40,):
            a = run(m, Q, T, thr, 0); t0 = clb.last_kernel_ms("match")
            for ver in VERS:
                clb.set_param("match_tc_ver", ver)
                b = run(m, Q, T, thr, 1); t1 = clb.last_kernel_ms("match")
                ok = all(np.array_equal(x, y) for x, y in zip(a, b))
                bad += not ok
                print(f"{'PASS' if ok else 'FAIL'} nQ={nQ} nT={nT} thr={thr} popc={t0:.3f} ms tc v{ver}={t1:.3f} ms "
                      f"({nQ * nT / max(t1, 1e-6) / 1e6:.1f} G cmp/s) matches={(a[0] >= 0).sum()}", flush=True)
                if not ok:
                    for name, x, y in zip(("match", "best", "second"), a, b):
                        d = np.nonzero(x != y)[0]
                        print(f"   {name}: {len(d)} differ; first {d[:5]} popc={x[d[:5]]} tc={y[d[:5]]}")
    # adversarial order: distances strictly decreasing along the DB (every row beats the previous best)
    nT = 4096
    q = rand_rows(rng, 1)
    bits = np.unpackbits(q.view(np.uint8), bitorder="little").reshape(1, 512)
    T = np.repeat(bits, nT, axis=0)
    for r in range(nT):
        k = 511 - (r * 511) // nT
        T[r, :k] ^= 1
    T = np.packbits(T, axis=1, bitorder="little").view(np.uint64)
    Q = np.repeat(q, 300, axis=0)
    a = run(m, Q, T, 0, 0); b = run(m, Q, T, 0, 1)
    ok = all(np.array_equal(x, y) for x, y in zip(a, b)); bad += not ok
    print(f"{'PASS' if ok else 'FAIL'} descending-distance DB", flush=True)
    # map path (expansion cached across calls)
    Q, T = planted(rng, 3000, 50000)
    clb.set_param("match_tc", 0); m.set_map_data(T); a = m.match_with_map(Q, 60, want_dist=True)
    clb.set_param("match_tc", 1); m.set_map_data(T); b = m.match_with_map(Q, 60, want_dist=True); c = m.match_with_map(Q[::-1].copy(), 60, want_dist=True)
    ok = all(np.array_equal(x, y) for x, y in zip(a, b)) and all(np.array_equal(x[::-1], y) for x, y in zip(a, c)); bad += not ok
    print(f"{'PASS' if ok else 'FAIL'} map path", flush=True)
    clb.set_param("match_tc", -1)
    print("tc_probe:", "ALL PASS" if bad == 0 else f"{bad} FAILED")
    sys.exit(1 if bad else 0)


if __name__ == "__main__":
    main()
