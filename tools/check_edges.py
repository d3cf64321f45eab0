"""Diagnostic: Canny / morphology maps of the CUDA image stage against the oracle on seeded cases,
with the location statistics of any mismatch (which rows, which columns modulo the 240-px strips).
  python tools/check_edges.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import mantis_b200  # noqa: E402
from cases import all_cases, blocky  # noqa: E402
from frames import K_for, poses  # noqa: E402
from oracle import pyoracle as po  # noqa: E402


def main():
    cases = dict(all_cases())
    rng = np.random.default_rng(5)
    cases["blocky_480x640"] = blocky(rng, 480, 640, 31)
    cases["noise_64x256"] = rng.integers(0, 256, (64, 256, 3), dtype=np.uint8)
    cases["noise_40x16"] = rng.integers(0, 256, (40, 16, 3), dtype=np.uint8)
    K, D = K_for(640, 480), np.zeros(4)
    for i, (R, t) in enumerate(poses(3, seed=77)):
        cases[f"grid_{i}"] = po.render_grid(R, t, K, D, 640, 480)
    bad = 0
    with mantis_b200.Context(0) as ctx:
        for name, img in cases.items():
            canny, morph = ctx.edges(img)
            oc, om = po.edge_stage(img, 50)
            d = np.argwhere(canny[0] != oc)
            msg = "ok"
            if len(d):
                bad += 1
                ys, xs = d[:, 0], d[:, 1]
                msg = (f"{len(d)} px differ; rows {ys.min()}..{ys.max()} cols {xs.min()}..{xs.max()}; "
                       f"x%240 hist {np.bincount(xs % 240 // 8, minlength=30).tolist()}; first {d[:6].tolist()}; "
                       f"gpu_only {int((canny[0][ys, xs] != 0).sum())} oracle_only {int((oc[ys, xs] != 0).sum())}")
            print(f"{name:>20} {img.shape}: {msg}", flush=True)
    print("MISMATCHING CASES:", bad)


if __name__ == "__main__":
    main()
