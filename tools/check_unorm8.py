"""Exhaustive check (exact rational arithmetic) of the division-free texel normalisation in csrc/pyramid.cu:
RN(v / 255) == fma(fma(-255, q, v), r, q) with q = RN(v * r), r = RN(1/255) = 0x3B808081, for v = 0..255."""
from fractions import Fraction
import numpy as np


def rn32(fr):
    """round a Fraction to the nearest float32, ties to even"""
    f = np.float32(float(fr))
    cands = [f, np.nextafter(f, np.float32(np.inf)), np.nextafter(f, np.float32(-np.inf))]
    return np.float32(min(cands, key=lambda c: (abs(Fraction(float(c)) - fr), int(np.float32(c).view(np.uint32)) & 1)))


def check():
    r = rn32(Fraction(1, 255))
    assert int(r.view(np.uint32)) == 0x3B808081
    plain_wrong = 0
    for v in range(256):
        true = rn32(Fraction(v, 255))
        q = rn32(Fraction(v) * Fraction(float(r)))
        plain_wrong += int(q != true)
        e = rn32(Fraction(v) - 255 * Fraction(float(q)))
        q2 = rn32(Fraction(float(e)) * Fraction(float(r)) + Fraction(float(q)))
        assert q2 == true, (v, true, q2)
    return plain_wrong


if __name__ == "__main__":
    print("corrected product exact for all 256 inputs; the plain product v * r is wrong for", check(), "of them")
