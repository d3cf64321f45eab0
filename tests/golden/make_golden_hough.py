"""Generates tests/golden/hough_golden.json: cv2.HoughLines (4.13.0) on the edge maps of the
shared image cases.  Row a6' has no reference call site (the reference has no Hough transform);
this pins the oracle's voting / peak stage (oracle/hough.c) to the OpenCV the reference links.

    python tests/golden/make_golden_hough.py     (build container only: needs cv2)
"""
import hashlib
import json
import os
import sys

import cv2
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
CASES = {"grid1": 120, "lena": 90, "blocky_720x1280": 150, "ramps_480x640": 100, "odd_33x129": 20, "noise_123x77": 40,
         "tiny_3x5": 1}


def edge_map(img, T=50):
    g = cv2.cvtColor(img, cv2.COLOR_BGR2GRAY)
    b = cv2.GaussianBlur(g, (3, 3), 3, 3)
    return cv2.Canny(b, T, 3 * T, apertureSize=3)


def digest(lines):
    a = np.ascontiguousarray(np.asarray(lines, np.float32).reshape(-1, 2))
    return hashlib.md5(a.tobytes()).hexdigest()


def main():
    from cases import all_cases
    cases = all_cases()
    out = {"cv2": cv2.__version__, "cases": {}}
    for name, thr in CASES.items():
        e = edge_map(cases[name])
        l = cv2.HoughLines(e, 1, np.pi / 180, thr)
        l = np.zeros((0, 2), np.float32) if l is None else l.reshape(-1, 2)
        out["cases"][name] = {"threshold": thr, "n_lines": int(len(l)), "lines_md5": digest(l),
                              "first": [[float(x) for x in r] for r in l[:4]]}
    with open(os.path.join(HERE, "hough_golden.json"), "w") as f:
        json.dump(out, f, indent=1)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
