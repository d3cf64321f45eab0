"""Records the oracle's whole-frame results (tests/golden/pipeline_golden.json) for grid1.jpg
and seeded synthetic frames.  The reference records no expected pose anywhere,
so these are the oracle's own outputs, generated in the build container right after
tests/test_oracle_parts.py / test_oracle_pipeline.py pinned its pieces against cv2 4.13 and
the compiled reference RPP.  Run:  python tests/golden/make_golden_pipeline.py"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from cases import all_cases  # noqa: E402
from frames import D_UNIT, K_GRID1, K_for, poses  # noqa: E402
from oracle import pyoracle as po  # noqa: E402


def record(tr, extra):
    d = dict(extra)
    d.update(status=tr.status, n_contours=tr.n_contours, n_quads_raw=tr.n_quads_raw, n_quads=tr.n_quads,
             n_hyps_generated=tr.n_hyps_generated, n_hyps_cluster=tr.n_hyps_cluster, n_scored=tr.n_scored)
    if tr.status == 0:
        d.update(pos=list(tr.pos), quat=list(tr.quat_xyzw), error=tr.error, min_yaw_diff=tr.min_yaw_diff,
                 first_best_error=tr.first_best.error, optimized_error=tr.optimized.error,
                 yaw_err=list(tr.yaw_err), best_rot=tr.best_rot)
    return d


def main():
    frames = []
    tr, _ = po.localize(all_cases()["grid1"], K_GRID1, D_UNIT)
    frames.append(record(tr, dict(source="grid1", K=K_GRID1.tolist(), D=D_UNIT.tolist(), w=907, h=505)))
    for (w, h, n, seed) in ((640, 480, 6, 0x4D414E54), (1280, 720, 2, 0x4D414E55)):
        K, D = K_for(w, h), np.zeros(4)
        for R, t in poses(n, seed=seed):
            img = po.render_grid(R, t, K, D, w, h)
            tr, _ = po.localize(img, K, D)
            frames.append(record(tr, dict(source="synthetic", K=K.tolist(), D=D.tolist(), w=w, h=h, R=R.tolist(),
                                          t=t.tolist())))
    json.dump({"frames": frames}, open(os.path.join(HERE, "pipeline_golden.json"), "w"), indent=1)
    for f in frames:
        print(f["source"], f["w"], f["status"], f["n_quads"], f.get("pos"), f.get("min_yaw_diff"))


if __name__ == "__main__":
    main()
