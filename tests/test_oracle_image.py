"""CPU: the C oracle's image stage against the cv2-generated golden digests (and against
cv2 itself when it is importable)."""
import hashlib
import json
import os

import numpy as np
import pytest

from cases import all_cases
from oracle import pyoracle as po

GOLD = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "image_stage_golden.json")))


def md5(a):
    return hashlib.md5(np.ascontiguousarray(a).tobytes()).hexdigest()


@pytest.mark.parametrize("name", sorted(GOLD["cases"]))
def test_oracle_matches_golden(name):
    img = all_cases()[name]
    g = GOLD["cases"][name]
    assert md5(img) == g["input_md5"]
    gray = po.bgr2gray(img)
    blur = po.blur3(gray)
    canny = po.canny(blur, 50, 150)
    morph = po.erode3(po.dilate3(canny, 2), 1)
    assert md5(gray) == g["gray_md5"]
    assert md5(blur) == g["blur_md5"]
    assert md5(canny) == g["canny_md5"]
    assert int((canny > 0).sum()) == g["edge_px"]
    assert md5(morph) == g["morph_md5"]
    c2, m2 = po.edge_stage(img, 50)
    assert md5(c2) == g["canny_md5"] and md5(m2) == g["morph_md5"]


def test_oracle_matches_cv2_live():
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(7)
    for shape in [(61, 93), (200, 333), (480, 640)]:
        base = rng.integers(0, 256, (shape[0] // 8 + 1, shape[1] // 8 + 1, 3), dtype=np.uint8)
        img = cv2.resize(base, (shape[1], shape[0]), interpolation=cv2.INTER_CUBIC)
        for T in (20, 50, 90):
            g = cv2.cvtColor(img, cv2.COLOR_BGR2GRAY)
            b = cv2.GaussianBlur(g, (3, 3), 3, 3)
            c = cv2.Canny(b, T, 3 * T, apertureSize=3)
            e = cv2.erode(cv2.dilate(c, None, iterations=2), None, iterations=1)
            oc, oe = po.edge_stage(img, T)
            assert np.array_equal(c, oc) and np.array_equal(e, oe)
