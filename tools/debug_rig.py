"""Diagnostic: rig-of-one (identity extrinsic) against the single-camera path, field by field."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import mantis_b200  # noqa: E402
from frames import K_for, poses  # noqa: E402
from oracle import pyoracle as po  # noqa: E402

K, D = K_for(640, 480), np.zeros(4)
frames = np.stack([po.render_grid(R, t, K, D, 640, 480) for R, t in poses(6, seed=0x4D414E54 + 61)])
with mantis_b200.Context(0) as ctx:
    s, ts = ctx.localize(frames, K, D, want_trace=True)
    r, tr = ctx.localize_rig(frames[:, None], K[None], D[None], np.array([[0, 0, 0, 0, 0, 0, 1.0]]), want_trace=True)
for i in range(len(frames)):
    o = po.localize_rig(frames[i][None], K[None], D[None], np.array([[0, 0, 0, 0, 0, 0, 1.0]]))
    print(i, "status", s[i].status, r[i].status, "nsel", ts[i].n_hyps_cluster, tr[i].n_hyps_cluster, "gen", ts[i].n_hyps_generated,
          tr[i].n_hyps_generated)
    print("   first_best", ts[i].first_best_error, tr[i].first_best_error, o.first_best.error)
    print("   optimized ", ts[i].optimized_error, tr[i].optimized_error, o.optimized.error)
    print("   shift_eq", list(ts[i].shift_err) == list(tr[i].shift_err), "yaw", list(ts[i].yaw_err), list(tr[i].yaw_err))
    print("   pos", list(s[i].pos), list(r[i].pos), "err", s[i].error, r[i].error, "npart", s[i].n_particles, r[i].n_particles)
i = 0
print("w2c single", [round(v, 6) for v in ts[i].optimized_w2c])
print("w2c rig   ", [round(v, 6) for v in tr[i].optimized_w2c])
print("w2c equal ", list(ts[i].optimized_w2c) == list(tr[i].optimized_w2c))
a, b = np.array(ts[i].shift_err), np.array(tr[i].shift_err)
print("shift single", a[:12])
print("shift rig   ", b[:12])
print("n differing", int((a != b).sum()), "sorted equal", bool(np.array_equal(np.sort(a), np.sort(b))))
w = np.array(ts[0].optimized_w2c)
gs, half = 0.32, 4.5 * 0.32
xs = []
x = -half + gs / 2
while x < half:
    xs.append(x); x += gs
cands = []
for x in xs:
    for y in xs:
        c = w.copy(); c[9] += x; c[10] += y
        cands.append(c)
cands = np.array(cands)
with mantis_b200.Context(0) as ctx:
    ext = ctx.score(frames[:1], K, D, cands[None])[0]
orc = np.array([po.score_white(c, frames[0], K, D) for c in cands[:12]])
print("ext   ", ext[:12])
print("oracle", orc)
print("ext==single", int((ext != a).sum()), "ext==rig", int((ext != b).sum()))
