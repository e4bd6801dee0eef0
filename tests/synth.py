"""Seeded synthetic inputs shared by tests, smoke() and bench.py."""
import numpy as np


def synth_image(H, W, seed, kind="U", black_frac=0.001):
    """U: i.i.d. uniform uint8 (worst case, ~all colours distinct).  N: natural-like, sum of
    three low-frequency cosines per channel + Gaussian noise sigma=6.  A fraction of pixels
    is forced to pure black to exercise the black-pixel exclusion (LST:134-135)."""
    rng = np.random.default_rng(seed)
    if kind == "U":
        img = rng.integers(0, 256, (H, W, 3), dtype=np.uint8)
    elif kind == "N":
        yy, xx = np.mgrid[0:H, 0:W].astype(np.float32)
        img = np.empty((H, W, 3), np.float32)
        for c in range(3):
            acc = np.zeros((H, W), np.float32)
            for _ in range(3):
                fx, fy = rng.uniform(0.5, 3.0, 2)
                ph = rng.uniform(0, 2 * np.pi)
                acc += np.cos(2 * np.pi * (fx * xx / W + fy * yy / H) + ph)
            img[..., c] = 128 + 40 * acc + rng.normal(0, 6, (H, W))
        img = np.clip(img, 0, 255).astype(np.uint8)
    else:
        raise ValueError(kind)
    if black_frac > 0:
        k = max(1, int(H * W * black_frac))
        idx = rng.choice(H * W, size=k, replace=False)
        img.reshape(-1, 3)[idx] = 0
    return img


def ellipse_mask(H, W, fg_frac=0.4, seed=4):
    """background_mask (1 = background) with an elliptical foreground of ~fg_frac of the
    pixels and a +-2 px jitter on the boundary."""
    rng = np.random.default_rng(seed)
    yy, xx = np.mgrid[0:H, 0:W]
    a = np.sqrt(fg_frac / np.pi)
    jit = rng.uniform(-2, 2, (H, W)) / max(H, W)
    inside = (((yy - H / 2) / (a * H)) ** 2 + ((xx - W / 2) / (a * W)) ** 2 + jit) <= 1.0
    return (~inside).astype(np.uint8)


def bernoulli_mask(H, W, p_bg=0.5, seed=4):
    rng = np.random.default_rng(seed)
    return (rng.random((H, W)) < p_bg).astype(np.uint8)


def voronoi_labels(H, W, n_labels=8, seed=5):
    rng = np.random.default_rng(seed)
    sites = rng.uniform(0, 1, (n_labels, 2)) * [H, W]
    yy, xx = np.mgrid[0:H, 0:W]
    d = (yy[None] - sites[:, 0, None, None]) ** 2 + (xx[None] - sites[:, 1, None, None]) ** 2
    return d.argmin(0).astype(np.uint8)
