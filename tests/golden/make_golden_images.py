"""Generates tests/golden/images.npz and image_stage_golden.json.

Run in the build container only (needs /root/reference and cv2 4.13.0 -- the OpenCV the
reference's cvtColor/GaussianBlur/Canny/dilate/erode calls resolve to here):
    python tests/golden/make_golden_images.py
The decoded fixture images are committed so that the GPU box (which has no
/root/reference) sees identical bytes; the digests pin the oracle and the CUDA path.
"""
import hashlib
import json
import os
import sys

import cv2
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference"


def md5(a):
    return hashlib.md5(np.ascontiguousarray(a).tobytes()).hexdigest()


def stages(img, T=50):
    g = cv2.cvtColor(img, cv2.COLOR_BGR2GRAY)           # QuadDetection.h:209
    b = cv2.GaussianBlur(g, (3, 3), 3, 3)                # :211
    c = cv2.Canny(b, T, 3 * T, apertureSize=3)           # :212
    d = cv2.dilate(c, None, iterations=2)                # :213
    e = cv2.erode(d, None, iterations=1)                 # :214
    return dict(gray=g, blur=b, canny=c, morph=e)


def main():
    grid1 = cv2.imread(os.path.join(REF, "test/grid1.jpg"))
    lena = cv2.imread(os.path.join(REF, "test/lena.jpg"))
    np.savez_compressed(os.path.join(HERE, "images.npz"), grid1=grid1, lena=lena)
    sys.path.insert(0, os.path.dirname(HERE))
    from cases import all_cases
    cases = all_cases()
    out = {"cv2": cv2.__version__, "cases": {}}
    for name, img in cases.items():
        s = stages(img)
        out["cases"][name] = {
            "shape": list(img.shape),
            "input_md5": md5(img),
            "edge_px": int((s["canny"] > 0).sum()),
            **{k + "_md5": md5(v) for k, v in s.items()},
        }
    with open(os.path.join(HERE, "image_stage_golden.json"), "w") as f:
        json.dump(out, f, indent=1)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
