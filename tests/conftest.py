import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

GOLDEN = os.path.join(ROOT, "tests", "golden")
PNGSUITE = os.path.join(GOLDEN, "pngsuite")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def oracle():
    import oracle_lib
    oracle_lib.build()
    oracle_lib.lib()
    return oracle_lib


def splitmix_rgba(w, h, seed=0x6E6F726D, alpha=None):
    """Deterministic test image: splitmix64(seed ^ pixel_index) low 32 bits as RGBA
    (SURVEY 8d config 3 generator), optionally with alpha forced."""
    idx = np.arange(w * h, dtype=np.uint64)
    z = (idx ^ np.uint64(seed)) + np.uint64(0x9E3779B97F4A7C15)
    z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
    z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
    z = z ^ (z >> np.uint64(31))
    px = (z & np.uint64(0xFFFFFFFF)).astype(np.uint32).view(np.uint8).reshape(h, w, 4).copy()
    if alpha is not None:
        px[..., 3] = alpha
    return px


def smooth_rgba(w, h, seed=1):
    """Band-limited colourful image (gradients + a few sinusoids) so blur / Sobel are non-trivial."""
    rng = np.random.default_rng(seed)
    y, x = np.mgrid[0:h, 0:w].astype(np.float64)
    out = np.zeros((h, w, 4), np.float64)
    for c in range(4):
        acc = rng.uniform(0.2, 0.8) + 0.3 * (x / max(w - 1, 1) - 0.5) * rng.uniform(-1, 1) \
            + 0.3 * (y / max(h - 1, 1) - 0.5) * rng.uniform(-1, 1)
        for _ in range(4):
            fx, fy, ph = rng.uniform(0.01, 0.3), rng.uniform(0.01, 0.3), rng.uniform(0, 6.28)
            acc = acc + 0.12 * np.sin(fx * x + fy * y + ph)
        out[..., c] = acc
    noise = rng.integers(-6, 7, size=out.shape)
    return np.clip(np.rint(out * 255 + noise), 0, 255).astype(np.uint8)
