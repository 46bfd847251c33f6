"""Deterministic synthetic images (the reference ships no sample; its README example is
``waterfall_1920_2520.raw``, README.md:22-23).

``waterfall_image``: smooth vertical streaks + low-frequency shading + noise — blurs the way a
photo does (values keep changing for hundreds of iterations).  ``random_image``: i.i.d. bytes,
the worst case for the convergence check and what the throughput benches use.
"""
from __future__ import annotations

import numpy as np


def random_image(height, width, channels=1, seed=0) -> np.ndarray:
    rng = np.random.default_rng(seed)
    shape = (height, width) if channels == 1 else (height, width, channels)
    return rng.integers(0, 256, size=shape, dtype=np.uint8)


def waterfall_image(height, width, channels=3, seed=0) -> np.ndarray:
    rng = np.random.default_rng(seed)
    y = np.linspace(0.0, 1.0, height, dtype=np.float32)[:, None]
    x = np.linspace(0.0, 1.0, width, dtype=np.float32)[None, :]
    streak = rng.random(width, dtype=np.float32)[None, :]
    streak = (streak + np.roll(streak, 1, 1) + np.roll(streak, 2, 1) + np.roll(streak, 3, 1)) / 4
    base = 0.55 * streak * (0.4 + 0.6 * y) + 0.25 * np.sin(6.0 * x + 3.0 * y) ** 2 + 0.2 * y
    planes = []
    for c in range(channels):
        noise = rng.normal(0.0, 0.04, size=(height, width)).astype(np.float32)
        planes.append(np.clip((base * (0.8 + 0.1 * c) + noise) * 255.0, 0, 255).astype(np.uint8))
    return planes[0] if channels == 1 else np.stack(planes, axis=2)
