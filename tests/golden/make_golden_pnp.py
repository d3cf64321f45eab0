"""Generates tests/golden/pnp_golden.npz: cv2.solvePnP(SOLVEPNP_ITERATIVE) (4.13.0) results for
 (a) the reference's own vectors, test/legacy/unit_test.cpp:49-72 (K = 300/150, unit square,
     rvec (-1, pi, 0), tvec (0, 0.6, 3)), and
 (b) 48 seeded config-5 cases: 200 of the 720 map points, pinhole 320/320/240, image points with
     N(0, 0.5 px) noise, float32 inputs (what the C ABI takes).
The OpenCV call is what the legacy path makes (include/legacy/mantis2/PoseEstimator.h:103); OpenCV is
third-party code absent from the reference tree, so these vectors are the pin for K11.

    python tests/golden/make_golden_pnp.py      (build container only: needs cv2)
"""
import os
import sys

import cv2
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))


def map_points():
    """params/map.yaml geometry: 0.08 m pitch points along the 10 + 10 grid lines (z = 0)."""
    pts = set()
    for k in range(10):
        c = round(-1.44 + 0.32 * k, 2)
        for j in range(37):
            v = round(-1.44 + 0.08 * j, 2)
            pts.add((c, v))
            pts.add((v, c))
    return np.array(sorted(pts), np.float64)


def main():
    from frames import poses
    K_leg = np.array([[300, 0, 150], [0, 300, 150], [0, 0, 1]], np.float64)
    sq = np.array([[.5, .5, 0], [-.5, .5, 0], [-.5, -.5, 0], [.5, -.5, 0]], np.float64)
    rvec, tvec = np.array([-1, np.pi, 0.]), np.array([0, 0.6, 3.])
    img, _ = cv2.projectPoints(sq, rvec, tvec, K_leg, None)
    img = img.reshape(-1, 2).astype(np.float32)
    ok, r, t = cv2.solvePnP(sq, img.astype(np.float64), K_leg, None, flags=cv2.SOLVEPNP_ITERATIVE)
    out = dict(legacy_obj=sq.astype(np.float32), legacy_img=img, legacy_K=K_leg.astype(np.float32),
               legacy_rvec=r.ravel(), legacy_tvec=t.ravel(), legacy_true_rvec=rvec, legacy_true_tvec=tvec)

    mp = map_points()
    mp3 = np.concatenate([mp, np.zeros((len(mp), 1))], 1)
    K5 = np.array([[320, 0, 320], [0, 320, 240], [0, 0, 1]], np.float64)
    rng = np.random.default_rng(0x4D414E54 + 5)
    objs, imgs, rs, ts = [], [], [], []
    for Rw, tw in poses(48, seed=0x4D414E54 + 5):
        Rc, tc = Rw.T, -Rw.T @ tw
        ob = mp3[rng.choice(len(mp3), 200, replace=False)].astype(np.float32)
        pc = (Rc @ ob.astype(np.float64).T).T + tc
        uv = (np.stack([320 * pc[:, 0] / pc[:, 2] + 320, 320 * pc[:, 1] / pc[:, 2] + 240], 1) +
              rng.normal(0, 0.5, (200, 2))).astype(np.float32)
        ok, r, t = cv2.solvePnP(ob.astype(np.float64), uv.astype(np.float64), K5, None, flags=cv2.SOLVEPNP_ITERATIVE)
        assert ok
        objs.append(ob); imgs.append(uv); rs.append(r.ravel()); ts.append(t.ravel())
    out.update(obj=np.stack(objs), img=np.stack(imgs), K=K5.astype(np.float32), rvec=np.stack(rs), tvec=np.stack(ts),
               cv2_version=np.array(cv2.__version__))
    np.savez_compressed(os.path.join(HERE, "pnp_golden.npz"), **out)
    print("wrote pnp_golden.npz", {k: np.asarray(v).shape for k, v in out.items()})


if __name__ == "__main__":
    main()
