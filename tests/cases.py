"""Seeded synthetic image cases shared by the golden generator and the tests (numpy only)."""
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SEED = 0x4D414E54  # "MANT"


def fixture_images():
    z = np.load(os.path.join(HERE, "golden", "images.npz"))
    return {"grid1": z["grid1"], "lena": z["lena"]}


def blocky(rng, h, w, cell):
    """Piecewise-constant random colour blocks with a little noise: long clean edges."""
    gh, gw = (h + cell - 1) // cell, (w + cell - 1) // cell
    base = rng.integers(0, 256, (gh, gw, 3), dtype=np.uint8)
    img = np.kron(base, np.ones((cell, cell, 1), dtype=np.uint8))[:h, :w]
    noise = rng.integers(-6, 7, (h, w, 3))
    return np.clip(img.astype(np.int32) + noise, 0, 255).astype(np.uint8)


def ramps(h, w):
    """Smooth gradients with weak-only edges chained to a few strong seeds."""
    y, x = np.mgrid[0:h, 0:w]
    img = np.zeros((h, w, 3), np.uint8)
    img[..., 0] = (x * 255 // max(w - 1, 1)).astype(np.uint8)
    img[..., 1] = ((x // 7 + y // 5) % 2 * 40 + 60).astype(np.uint8)
    img[..., 2] = ((y * 3) % 256).astype(np.uint8)
    img[h // 3: h // 3 + 2, :, :] = 255
    return img


def synthetic_cases():
    rng = np.random.default_rng(SEED)
    return {
        "noise_123x77": rng.integers(0, 256, (123, 77, 3), dtype=np.uint8),
        "noise_480x640": rng.integers(0, 256, (480, 640, 3), dtype=np.uint8),
        "blocky_720x1280": blocky(rng, 720, 1280, 37),
        "blocky_1080x1920": blocky(rng, 1080, 1920, 53),
        "ramps_480x640": ramps(480, 640),
        "tiny_3x5": rng.integers(0, 256, (3, 5, 3), dtype=np.uint8),
        "row_1x64": rng.integers(0, 256, (1, 64, 3), dtype=np.uint8),
        "col_70x1": rng.integers(0, 256, (70, 1, 3), dtype=np.uint8),
        "odd_33x129": blocky(rng, 33, 129, 9),
    }


def all_cases():
    c = dict(fixture_images())
    c.update(synthetic_cases())
    return c
