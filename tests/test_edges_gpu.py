"""GPU: the CUDA image stage (through the C ABI) against the oracle and the golden digests.
Bit-exact: Canny map and the dilate x2 / erode x1 map."""
import hashlib
import json
import os

import numpy as np
import pytest

from cases import all_cases, blocky
from oracle import pyoracle as po

pytestmark = pytest.mark.gpu
GOLD = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "image_stage_golden.json")))


def md5(a):
    return hashlib.md5(np.ascontiguousarray(a).tobytes()).hexdigest()


@pytest.mark.parametrize("name", sorted(GOLD["cases"]))
def test_edges_match_golden_and_oracle(ctx, name):
    img = all_cases()[name]
    canny, morph = ctx.edges(img)
    oc, om = po.edge_stage(img, 50)
    assert np.array_equal(canny[0], oc), f"canny differs at {np.argwhere(canny[0] != oc)[:5]}"
    assert np.array_equal(morph[0], om), f"morph differs at {np.argwhere(morph[0] != om)[:5]}"
    assert md5(canny[0]) == GOLD["cases"][name]["canny_md5"]
    assert md5(morph[0]) == GOLD["cases"][name]["morph_md5"]


def test_edges_batch_mixed(ctx):
    rng = np.random.default_rng(11)
    frames = np.stack([blocky(rng, 480, 640, c) for c in (11, 23, 47, 64, 90, 5)])
    canny, morph = ctx.edges(frames)
    for i in range(len(frames)):
        oc, om = po.edge_stage(frames[i], 50)
        assert np.array_equal(canny[i], oc)
        assert np.array_equal(morph[i], om)


def test_edges_long_weak_chain(ctx):
    """A 1-px strong seed at the end of a long weak snake: closure must travel the whole way."""
    h, w = 480, 640
    img = np.full((h, w, 3), 20, np.uint8)
    # faint serpentine (weak-only contrast) ...
    for k, y in enumerate(range(20, 460, 40)):
        img[y:y + 2, 10:630] = 50
        x = 626 if k % 2 == 0 else 10
        img[y:y + 42, x:x + 2] = 50
    img[20:22, 10:14] = 255  # ... with one strong seed
    canny, morph = ctx.edges(img)
    oc, om = po.edge_stage(img, 50)
    assert oc.sum() > 0
    assert np.array_equal(canny[0], oc)
    assert np.array_equal(morph[0], om)


def test_edges_idempotent_and_deterministic(ctx):
    img = all_cases()["grid1"]
    a = ctx.edges(img)
    b = ctx.edges(img)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    # the edge map only takes the values 0 / 255
    assert set(np.unique(a[0])) <= {0, 255}
