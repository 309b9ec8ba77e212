"""Seeded synthetic image generator for tests and bench: low-pass noise + fine noise + gradient +
hard-edged rectangles, which lands near photographic bit rates (≈1.5-2.5 bpp at q90).  Pure numpy; the JPEG bytes
themselves are produced by libjpeg-turbo (PIL's bundled copy in bench.py, oracle/_ref's in tests)."""
import numpy as np


try:
    import cv2 as _cv2
except Exception:  # pragma: no cover
    _cv2 = None


def _lowpass(rng, h, w, cell):
    """cheap band-limited noise: bilinear upsampling of a coarse random grid"""
    gh, gw = max(2, h // cell + 2), max(2, w // cell + 2)
    g = rng.standard_normal((gh, gw)).astype(np.float32)
    if _cv2 is not None and h * w > 65536:
        return _cv2.resize(g, (w, h), interpolation=_cv2.INTER_LINEAR)
    ys = np.linspace(0, gh - 1.001, h, dtype=np.float32)
    xs = np.linspace(0, gw - 1.001, w, dtype=np.float32)
    y0 = ys.astype(np.int32)
    x0 = xs.astype(np.int32)
    fy = (ys - y0)[:, None]
    fx = (xs - x0)[None, :]
    a = g[y0][:, x0]
    b = g[y0][:, x0 + 1]
    c = g[y0 + 1][:, x0]
    d = g[y0 + 1][:, x0 + 1]
    return a * (1 - fy) * (1 - fx) + b * (1 - fy) * fx + c * fy * (1 - fx) + d * fy * fx


def synth_image(seed, width, height, channels=3, kind="photo"):
    """-> (H, W, 3) or (H, W) uint8.  kind: photo | flat | noise"""
    rng = np.random.Generator(np.random.PCG64(seed))
    shape = (height, width, channels) if channels == 3 else (height, width)
    if kind == "flat":
        return np.full(shape, 137, np.uint8)
    if kind == "noise":
        return rng.integers(0, 256, shape, dtype=np.uint8)
    base = _lowpass(rng, height, width, max(4, width // 16))
    yy, xx = np.mgrid[0:height, 0:width].astype(np.float32)
    grad = (xx / max(1, width) - 0.5) * 40 + (yy / max(1, height) - 0.5) * 25
    chans = []
    for c in range(channels if channels == 3 else 1):
        own = _lowpass(rng, height, width, max(4, width // 16))
        v = 128 + 80 * (0.7 * base + 0.3 * own) + 30 * _lowpass(rng, height, width, 4) \
            + 6 * rng.standard_normal((height, width)).astype(np.float32) + grad
        chans.append(v)
    img = np.stack(chans, -1) if channels == 3 else chans[0]
    for _ in range(8):
        x0, y0 = int(rng.integers(0, width)), int(rng.integers(0, height))
        x1 = min(width, x0 + int(rng.integers(1, max(2, width // 4))))
        y1 = min(height, y0 + int(rng.integers(1, max(2, height // 4))))
        img[y0:y1, x0:x1] = rng.integers(0, 256, (channels,) if channels == 3 else ()).astype(np.float32)
    return np.clip(img, 0, 255).astype(np.uint8)
