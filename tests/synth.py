"""Seeded synthetic inputs shared by the tests, bench.py and smoke().

Frames: uniform uint8 noise -> cv2/numpy Gaussian (sigma 2) -> min-max normalised to 0..255: dense
corners on every pyramid level.  `contrast` < 1 squeezes the grey range around 128 so that the
minThFAST fallback of the per-cell FAST (reference src/ORBextractor.cc:764-770) is exercised.
Pure numpy on purpose (no cv2) so the generator is identical wherever it runs.
"""
import numpy as np


def _gauss_kernel(sigma=2.0, radius=6):
    x = np.arange(-radius, radius + 1, dtype=np.float64)
    k = np.exp(-(x * x) / (2 * sigma * sigma))
    return k / k.sum()


def _blur(img):
    k = _gauss_kernel()
    r = len(k) // 2
    p = np.pad(img, ((0, 0), (r, r)), mode="reflect")
    out = np.zeros_like(img)
    for i, kv in enumerate(k):
        out += kv * p[:, i:i + img.shape[1]]
    p = np.pad(out, ((r, r), (0, 0)), mode="reflect")
    out2 = np.zeros_like(img)
    for i, kv in enumerate(k):
        out2 += kv * p[i:i + img.shape[0], :]
    return out2


def textured_frame(w, h, seed, contrast=1.0):
    rng = np.random.default_rng(seed)
    a = rng.integers(0, 256, size=(h, w)).astype(np.float64)
    a = _blur(a)
    a = (a - a.min()) / (a.max() - a.min())
    if contrast != 1.0:
        a = 0.5 + (a - 0.5) * contrast
    return np.clip(np.rint(a * 255.0), 0, 255).astype(np.uint8)


def mixed_frame(w, h, seed):
    """Left half full contrast, right half low contrast (fallback cells), a flat band (empty cells)."""
    a = textured_frame(w, h, seed).astype(np.float64)
    b = 128 + (a - 128) * 0.08
    out = a.copy()
    out[:, w // 2:] = b[:, w // 2:]
    out[h // 3: h // 3 + 45, :] = 97
    return np.clip(np.rint(out), 0, 255).astype(np.uint8)


def frame_batch(w, h, n, seed0=0, distinct=8):
    """n frames built from `distinct` seeded textures (rolled so that every frame differs)."""
    base = [textured_frame(w, h, seed0 + i) for i in range(min(distinct, n))]
    out = np.empty((n, h, w), np.uint8)
    for i in range(n):
        out[i] = np.roll(base[i % len(base)], shift=(3 * (i // len(base)), 5 * (i // len(base))), axis=(0, 1))
    return out


def random_descriptors(n, seed):
    rng = np.random.default_rng(seed)
    return rng.integers(0, 256, size=(n, 32), dtype=np.uint8)


def flip_bits(desc, nflips, rng):
    """Return a copy of desc (n,32) with nflips[i] random distinct bit flips in row i."""
    out = desc.copy()
    for i in range(desc.shape[0]):
        k = int(nflips[i])
        if k <= 0:
            continue
        bits = rng.choice(256, size=k, replace=False)
        for b in bits:
            out[i, b >> 3] ^= np.uint8(1 << (b & 7))
    return out
