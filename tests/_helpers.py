"""test-only helpers (checkers, never product code)."""
import numpy as np


def merge_partials_host(parts, threshold):
    """numpy statement of the merge rule, for host-side checks of gathered partials (test-side checker; the product merges
    on the device, clb_merge_partials_dev).  parts: [P, nQ] PARTIAL_DTYPE in ascending
    DB order.  Returns (match, best, second)."""
    parts = np.asarray(parts)
    nq = parts.shape[1]
    best = np.full(nq, 100000, np.int32); second = np.full(nq, 200000, np.int32); bi = np.full(nq, -1, np.int32)
    any_ = np.zeros(nq, bool)
    for p in parts:
        ok = p["best_i"] >= 0
        any_ |= ok
        for d, idx in ((p["best_d"], p["best_i"]), (p["second_d"], None)):
            d = np.where(ok, d, np.int32(2 ** 30))
            second = np.minimum(second, d)
            lt = d < best
            second = np.where(lt, best, second)
            if idx is not None:
                bi = np.where(lt, idx, bi)
            best = np.where(lt, d, best)
    match = np.where(any_ & (second - best > threshold), bi, -1).astype(np.int32)
    return match, best, second




import functools


@functools.lru_cache(maxsize=4)
def filter_problem(seed=3):
    """two descriptor sets for the symmetric / ratio filter tests: noisy copies at several noise levels (one A row with several B
    candidates), near-duplicate A rows (two A rows sharing their nearest B row: the cross-check must drop one), A rows without a
    counterpart, B rows without one.  Nearest rows are unique (no distance ties at the minimum), so openMVG's unspecified tie
    order cannot show."""
    from coloc_b200 import synth
    rng = np.random.default_rng(seed)

    def noisy(rows, p, sd):
        r = np.random.default_rng(sd)
        bits = r.random((len(rows), 8, 64)) < p
        return rows ^ (bits.astype(np.uint64) << np.arange(64, dtype=np.uint64)).sum(axis=2, dtype=np.uint64)

    A = synth.random_descriptors(400, seed)
    A[200:250] = noisy(A[150:200], 0.03, seed + 1)                      # near-duplicates inside A
    B = np.concatenate([noisy(A[:300], 0.08, seed + 2), noisy(A[:120], 0.15, seed + 3), noisy(A[350:], 0.05, seed + 4),
                        synth.random_descriptors(60, seed + 5)])
    B = B[rng.permutation(len(B))]
    # remove distance ties at the minimum in both directions (openMVG's partial_sort picks an unspecified row among equals;
    # CUDAK2NN.cu:67-72 the first): pull one of the tied rows a bit closer until every nearest row is unique
    bitsA = np.unpackbits(A.view(np.uint8), axis=1).astype(np.int16)
    for _ in range(50):
        bitsB = np.unpackbits(B.view(np.uint8), axis=1).astype(np.int16)
        D = (bitsA[:, None, :] != bitsB[None, :, :]).sum(axis=2)
        changed = False
        for Dm, flip_b_of in ((D, lambda i, j: (i, j)), (D.T, lambda j, i: (i, j))):
            part = np.partition(Dm, 1, axis=1)
            for r in np.nonzero(part[:, 0] == part[:, 1])[0]:
                c = int(np.argmin(Dm[r]))
                i, j = flip_b_of(int(r), c)
                diff = np.nonzero(bitsA[i] != np.unpackbits(B[j].view(np.uint8)).astype(np.int16))[0]
                if len(diff):
                    byte, bit = divmod(int(diff[0]), 8)
                    Bb = B.view(np.uint8).reshape(len(B), 64)
                    Bb[j, byte] ^= np.uint8(0x80 >> bit)
                    changed = True
        if not changed:
            break
    return A, B
