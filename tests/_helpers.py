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


