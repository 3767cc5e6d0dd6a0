"""Synthetic inputs for parity tests and bench: textured frames with known
random homographies, and random / planted 512-bit descriptor sets.  Pure numpy, deterministic
per seed; no reference or oracle code involved."""
import numpy as np


def _value_noise(rng, h, w, octaves=6):
    img = np.zeros((h, w), np.float32)
    amp, tot = 1.0, 0.0
    for o in range(octaves):
        cell = max(2, 256 >> o)
        gh, gw = h // cell + 3, w // cell + 3
        g = rng.random((gh, gw), dtype=np.float32)
        ys = np.arange(h, dtype=np.float32) / cell
        xs = np.arange(w, dtype=np.float32) / cell
        y0 = ys.astype(np.int32); x0 = xs.astype(np.int32)
        fy = (ys - y0)[:, None]; fx = (xs - x0)[None, :]
        fy = fy * fy * (3 - 2 * fy); fx = fx * fx * (3 - 2 * fx)
        a = g[y0][:, x0]; b = g[y0][:, x0 + 1]; c = g[y0 + 1][:, x0]; d = g[y0 + 1][:, x0 + 1]
        img += amp * ((a * (1 - fx) + b * fx) * (1 - fy) + (c * (1 - fx) + d * fx) * fy)
        tot += amp; amp *= 0.6
    return img / tot


def textured_frame(w, h, seed, shapes=None, max_size=None):
    """uint8 frame: multi-octave value noise + random filled rectangles and discs (corners for FAST).
    `shapes` / `max_size` (pixels) set the corner density: the defaults give ~9k keypoints at 1280x720 (threshold 40),
    bench.py passes denser settings to reach the keypoint counts BASELINE.json's configs name."""
    rng = np.random.default_rng(seed)
    img = _value_noise(rng, h, w) * 255.0
    n = shapes if shapes is not None else max(40, (w * h) // 600)
    yy, xx = None, None
    for _ in range(n):
        val = float(rng.integers(0, 256))
        if rng.random() < 0.6:
            x0 = int(rng.integers(0, w)); y0 = int(rng.integers(0, h))
            rw = int(rng.integers(4, max(5, max_size or w // 12))); rh = int(rng.integers(4, max(5, max_size or h // 12)))
            img[y0:y0 + rh, x0:x0 + rw] = 0.35 * img[y0:y0 + rh, x0:x0 + rw] + 0.65 * val
        else:
            cx = int(rng.integers(0, w)); cy = int(rng.integers(0, h)); r = int(rng.integers(3, max(4, (max_size // 2) if max_size else h // 20)))
            ya, yb = max(0, cy - r), min(h, cy + r + 1); xa, xb = max(0, cx - r), min(w, cx + r + 1)
            yy, xx = np.mgrid[ya:yb, xa:xb]
            m = (yy - cy) ** 2 + (xx - cx) ** 2 <= r * r
            sub = img[ya:yb, xa:xb]
            sub[m] = 0.35 * sub[m] + 0.65 * val
    return np.clip(img + 0.5, 0, 255).astype(np.uint8)


# corner densities that reach the keypoint counts BASELINE.json's configs name at FAST threshold 40
# (C1 640x480 ~5k, C2/C3 1280x720 ~20k, C5 3840x2160 ~50k); (shapes, max_size) for textured_frame
CONFIG_DENSITY = {(640, 480): (650, 40), (1280, 720): (3900, 48), (3840, 2160): (4900, 64)}


def config_frame(w, h, k=0):
    """frame k of a BASELINE config: seed 1000+k, density from CONFIG_DENSITY (default density for other sizes)."""
    sh, ms = CONFIG_DENSITY.get((w, h), (None, None))
    return textured_frame(w, h, 1000 + k, sh, ms)


def config_frame_pair(w, h, k=0):
    a = config_frame(w, h, k)
    H = random_homography(w, h, 2000 + k)
    return a, warp_homography(a, H), H


def config_views(w, h, n, k=0):
    """n views of the same scene (config 3: n vehicles looking at one area): view 0 is config_frame(w, h, k), view i > 0 is
    its warp by the random homography of seed 2000 + 16 k + i.  Returns (views, homographies) with H[0] = identity."""
    base = config_frame(w, h, k)
    views, Hs = [base], [np.eye(3)]
    for i in range(1, n):
        H = random_homography(w, h, 2000 + 16 * k + i)
        views.append(warp_homography(base, H)); Hs.append(H)
    return views, Hs


def random_homography(w, h, seed):
    """H = K R(rx,ry,rz) K^-1 composed with scale + shift; angles U(-10,10) deg, scale U(.8,1.25), shift U(-5%,5%)."""
    rng = np.random.default_rng(seed)
    rx, ry = np.deg2rad(rng.uniform(-10, 10, 2)) * 0.15   # out-of-plane kept mild so the frame stays covered
    rz = np.deg2rad(rng.uniform(-10, 10))
    s = rng.uniform(0.8, 1.25)
    tx, ty = rng.uniform(-0.05, 0.05, 2) * (w, h)
    f = 0.9 * w
    K = np.array([[f, 0, w / 2], [0, f, h / 2], [0, 0, 1.0]])
    cx, sx = np.cos(rx), np.sin(rx); cy, sy = np.cos(ry), np.sin(ry); cz, sz = np.cos(rz), np.sin(rz)
    R = (np.array([[cz, -sz, 0], [sz, cz, 0], [0, 0, 1]]) @ np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
         @ np.array([[1, 0, 0], [0, cx, -sx], [0, sx, cx]]))
    S = np.array([[s, 0, (1 - s) * w / 2 + tx], [0, s, (1 - s) * h / 2 + ty], [0, 0, 1.0]])
    H = S @ K @ R @ np.linalg.inv(K)
    return H / H[2, 2]


def warp_homography(img, H):
    """second view: out(p) = bilinear img(H^-1 p), edge-clamped.  Done once on the host."""
    h, w = img.shape
    Hi = np.linalg.inv(H)
    ys, xs = np.mgrid[0:h, 0:w].astype(np.float64)
    d = Hi[2, 0] * xs + Hi[2, 1] * ys + Hi[2, 2]
    u = (Hi[0, 0] * xs + Hi[0, 1] * ys + Hi[0, 2]) / d
    v = (Hi[1, 0] * xs + Hi[1, 1] * ys + Hi[1, 2]) / d
    u = np.clip(u, 0, w - 1.001); v = np.clip(v, 0, h - 1.001)
    x0 = u.astype(np.int32); y0 = v.astype(np.int32)
    fx = (u - x0).astype(np.float32); fy = (v - y0).astype(np.float32)
    f = img.astype(np.float32)
    out = (f[y0, x0] * (1 - fx) + f[y0, x0 + 1] * fx) * (1 - fy) + (f[y0 + 1, x0] * (1 - fx) + f[y0 + 1, x0 + 1] * fx) * fy
    return np.clip(out + 0.5, 0, 255).astype(np.uint8)


def frame_pair(w, h, k=0):
    """frame k (seed 1000+k) and its homography-warped second view (seed 2000+k).  Returns (a, b, H)."""
    a = textured_frame(w, h, 1000 + k)
    H = random_homography(w, h, 2000 + k)
    return a, warp_homography(a, H), H


def random_descriptors(n, seed):
    """n x 8 uint64, i.i.d. uniform bits."""
    rng = np.random.default_rng(seed)
    return rng.integers(0, 2 ** 64, size=(n, 8), dtype=np.uint64)


def planted_queries(db, n, seed, flip_p=0.10, frac_true=0.8):
    """queries: frac_true are DB rows with each bit flipped w.p. flip_p (true matches), rest fresh random.
    Returns (Q, src) with src[i] = DB row index or -1."""
    rng = np.random.default_rng(seed)
    nt = int(n * frac_true)
    src = np.full(n, -1, np.int64)
    src[:nt] = rng.integers(0, len(db), nt)
    Q = rng.integers(0, 2 ** 64, size=(n, 8), dtype=np.uint64)
    # Bernoulli(flip_p) bit masks, 64 bits at a time
    bits = rng.random((nt, 8, 64), dtype=np.float32) < flip_p
    mask = (bits.astype(np.uint64) << np.arange(64, dtype=np.uint64)).sum(axis=2, dtype=np.uint64)
    Q[:nt] = db[src[:nt]] ^ mask
    perm = rng.permutation(n)
    return Q[perm], src[perm]
