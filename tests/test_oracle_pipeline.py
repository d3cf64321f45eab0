"""CPU: the oracle's whole-frame restatement (src/mantis3.cpp:68-135) -- quad extraction and
scoring cross-checked against the same OpenCV calls through cv2, end-to-end sanity on
synthetic frames, and the committed golden traces."""
import json
import os

import numpy as np
import pytest

from cases import all_cases
from frames import D_UNIT, K_GRID1, K_for, poses, rig_cam_to_base, rig_extrinsics
from oracle import pyoracle as po

HERE = os.path.dirname(os.path.abspath(__file__))


def quads_via_cv2(cv2, img, T=50, eps=10, mult=0.1, flann=False):
    mono = cv2.GaussianBlur(cv2.cvtColor(img, cv2.COLOR_BGR2GRAY), (3, 3), 3, 3)
    canny = cv2.Canny(mono, T, 3 * T, apertureSize=3)
    canny = cv2.erode(cv2.dilate(canny, None, iterations=2), None, iterations=1)
    cs, _ = cv2.findContours(canny, cv2.RETR_CCOMP, cv2.CHAIN_APPROX_SIMPLE)
    quads = []
    for c in cs:
        a = cv2.approxPolyDP(c, eps, True).reshape(-1, 2)
        if len(a) == 4:
            quads.append(a)
    n_raw = len(quads)
    # Quadrilateral ctor in float32 + removeDuplicateQuads with exact radius search
    f32 = np.float32
    cen = []
    for q in quads:
        xs, ys = f32(0), f32(0)
        for p in q:
            xs, ys = f32(xs + f32(p[0])), f32(ys + f32(p[1]))
        cen.append((f32(xs / f32(4)), f32(ys / f32(4))))
    side = [np.sqrt(float((q[0][0] - q[1][0]) ** 2 + (q[0][1] - q[1][1]) ** 2)) for q in quads]
    neighbor = [False] * n_raw
    tree = None
    if flann and n_raw:
        # the reference's own call (QuadDetection.h:130,143): KD-tree with one tree, radiusSearch(query, indices, dists,
        # radius, 4).  Its fresh std::vector<int> is resized to 4 zero-initialised entries by OpenCV, which the
        # zero-filled output arrays reproduce; FLANN writes only the hits.
        feats = np.array(cen, np.float32).reshape(-1, 2)
        tree = cv2.flann_Index(feats, dict(algorithm=1, trees=1))
    for e in range(n_raw):
        if neighbor[e]:
            continue
        rad = f32(mult * side[e])
        if tree is not None:
            ind, dst = np.zeros((1, 4), np.int32), np.zeros((1, 4), np.float32)
            tree.radiusSearch(np.array([cen[e]], np.float32), float(mult * side[e]), 4, ind, dst)
            for j in ind[0, 1:]:
                neighbor[int(j)] = True
            continue
        d = [f32(f32((cen[e][0] - c[0]) * (cen[e][0] - c[0])) + f32((cen[e][1] - c[1]) * (cen[e][1] - c[1]))) for c in cen]
        hits = sorted([(d[j], j) for j in range(n_raw) if d[j] <= rad])[:4]
        idx = [j for _, j in hits] + [0] * (4 - len(hits))
        for j in idx[1:]:
            neighbor[j] = True
    keep = [q for q, nb in zip(quads, neighbor) if not nb]
    return n_raw, keep, len(cs)


@pytest.mark.parametrize("name", ["grid1", "lena", "blocky_720x1280"])
def test_quads_match_cv2(name):
    cv2 = pytest.importorskip("cv2")
    img = all_cases()[name]
    n_raw, keep, n_cont = quads_via_cv2(cv2, img)
    tr, _ = po.localize(img, K_GRID1, D_UNIT)
    assert tr.n_contours == n_cont and tr.n_quads_raw == n_raw and tr.n_quads == len(keep)
    for i, q in enumerate(keep):
        assert list(tr.quad_corners[i]) == q.reshape(-1).tolist()
    if name == "grid1":
        assert (n_cont, n_raw) == (142, 24)  # SURVEY section 4 golden counts


@pytest.mark.parametrize("name", ["grid1", "lena", "blocky_720x1280", "blocky_1080x1920", "grid_a", "grid_b", "grid_c", "grid_d"])
def test_duplicate_removal_matches_cv2_flann(name):
    """removeDuplicateQuads through the real cv::flann::Index (KDTreeIndexParams(1), radiusSearch, maxResults 4):
    which quads survive -- hit order, the squared-radius comparison and the 'unfilled slots are index 0' quirk
    included -- must be what the oracle's exact search keeps."""
    cv2 = pytest.importorskip("cv2")
    if name.startswith("grid_"):
        K, D = K_for(640, 480), np.zeros(4)
        R, t = poses(4, seed=0x4D414E54 + 31)["abcd".index(name[-1])]
        img = po.render_grid(R, t, K, D, 640, 480)
        tr, _ = po.localize(img, K, D)
    else:
        img = all_cases()[name]
        tr, _ = po.localize(img, K_GRID1, D_UNIT)
    n_raw, keep, n_cont = quads_via_cv2(cv2, img, flann=True)
    assert (tr.n_contours, tr.n_quads_raw) == (n_cont, n_raw)
    assert tr.n_quads == len(keep)
    for i, q in enumerate(keep):
        assert list(tr.quad_corners[i]) == q.reshape(-1).tolist()


def score_white_numpy(cv2, hyp_c2w_R, hyp_c2w_t, img, mask, K, D, maps):
    pts = np.concatenate(maps)
    pc = pts @ hyp_c2w_R.T + hyp_c2w_t
    pc = pc[pc[:, 2] > 0]
    nrm = pc[:, :2] / pc[:, 2:3]
    px = cv2.fisheye.distortPoints(nrm.reshape(-1, 1, 2), K, D).reshape(-1, 2)
    h, w = mask.shape
    ok = (px[:, 0] < w) & (px[:, 1] < h) & (px[:, 0] >= 0) & (px[:, 1] > 0)
    px = px[ok]
    xi, yi = np.rint(px[:, 0]).astype(int), np.rint(px[:, 1]).astype(int)
    cleaned = np.where(mask[..., None] > 0, img, 0).reshape(-1, 3).astype(np.int64)
    lin = yi * w + xi
    inside = lin < h * w
    vals = np.zeros((len(lin), 3), np.int64)
    vals[inside] = cleaned[lin[inside]]
    err = ((vals - 255) ** 2).sum()
    return err / (len(px) * 1.1) if len(px) else np.finfo(np.float64).max


def test_first_evaluation_matches_numpy_cv2():
    cv2 = pytest.importorskip("cv2")
    K, D = K_for(640, 480), np.zeros(4)
    maps = po.make_map()
    for R, t in poses(4, seed=21):
        img = po.render_grid(R, t, K, D, 640, 480)
        tr, mask = po.localize(img, K, D, want_mask=True)
        if tr.status != 0:
            continue
        fb = tr.first_best
        e = score_white_numpy(cv2, np.array(fb.c2w.R[:]).reshape(3, 3), np.array(fb.c2w.t[:]), img, mask, K, D, maps)
        assert e == fb.error


def test_synthetic_localization_sane():
    """A full-grid view localizes to the ground-truth cell (the reference algorithm is not
    always right; these seeded poses are ones where it is)."""
    K, D = K_for(640, 480), np.zeros(4)
    good = 0
    for R, t in poses(6, seed=0x4D414E54):
        img = po.render_grid(R, t, K, D, 640, 480)
        tr, _ = po.localize(img, K, D)
        assert tr.status == 0
        if np.linalg.norm(np.array(tr.pos[:]) - t) < 0.05:
            good += 1
    assert good >= 1


def test_rng_is_reseeded_per_frame():
    K, D = K_for(640, 480), np.zeros(4)
    R, t = poses(1, seed=3)[0]
    img = po.render_grid(R, t, K, D, 640, 480)
    a, _ = po.localize(img, K, D)
    b, _ = po.localize(img, K, D)
    assert a.pos[:] == b.pos[:] and a.error == b.error


def test_edge_cases():
    K, D = K_for(64, 48), np.zeros(4)
    blank = np.zeros((48, 64, 3), np.uint8)
    tr, _ = po.localize(blank, K, D)
    assert tr.status == 1  # no quads -> early return, src/mantis3.cpp:83-84


def test_pipeline_golden():
    path = os.path.join(HERE, "golden", "pipeline_golden.json")
    gold = json.load(open(path))
    for g in gold["frames"]:
        K, D = np.array(g["K"], np.float32), np.array(g["D"])
        if g["source"] == "grid1":
            img = all_cases()["grid1"]
        else:
            img = po.render_grid(np.array(g["R"]), np.array(g["t"]), K, D, g["w"], g["h"])
        tr, _ = po.localize(img, K, D)
        assert tr.status == g["status"] and tr.n_quads == g["n_quads"] and tr.n_hyps_cluster == g["n_hyps_cluster"]
        if tr.status == 0:
            assert np.allclose(tr.pos[:], g["pos"], atol=1e-9) and np.allclose(tr.quat_xyzw[:], g["quat"], atol=1e-9)
            assert tr.error == pytest.approx(g["error"], rel=1e-12)


def _same_trace(a, b):
    return (a.status == b.status and a.n_quads == b.n_quads and a.n_hyps_generated == b.n_hyps_generated
            and a.n_hyps_cluster == b.n_hyps_cluster and a.first_best.error == b.first_best.error
            and a.optimized.error == b.optimized.error and list(a.shift_err) == list(b.shift_err)
            and list(a.yaw_err) == list(b.yaw_err) and list(a.pos) == list(b.pos) and list(a.quat_xyzw) == list(b.quat_xyzw)
            and a.error == b.error and a.n_scored == b.n_scored)


def test_rig_of_one_camera_is_the_single_camera_path():
    """mo_localize_rig with one camera at the base origin = quadDetection() itself, bit for bit."""
    K, D = K_for(640, 480), np.zeros(4)
    R, t = poses(1, seed=0x4D414E54 + 51)[0]
    img = po.render_grid(R, t, K, D, 640, 480)
    single, _ = po.localize(img, K, D)
    assert single.status == 0
    assert _same_trace(po.localize_rig(img[None], K[None], D[None], None), single)
    assert _same_trace(po.localize_rig(img[None], K[None], D[None], np.array([[0, 0, 0, 0, 0, 0, 1.0]])), single)


def test_rig_joint_localization_recovers_the_base_pose():
    """Two cameras of SURVEY 8d's rig (down + 30 degrees fore): hypotheses of both cameras pooled in the base frame,
    joint score = summed error / summed projected points (Mantis2Error.h:55-99).  The base pose comes back to the
    accuracy the stochastic refinement has (a few cm / a few 1e-2 rad, modulo the grid's symmetries)."""
    K, D = K_for(640, 480), np.zeros(4)
    ext = rig_extrinsics()[:2]
    Rb, tb = poses(1, seed=0x4D414E54 + 52)[0]
    frames = np.stack([po.render_grid(Rb @ Rc, tb + Rb @ tc, K, D, 640, 480) for Rc, tc in ext])
    tr = po.localize_rig(frames, np.stack([K] * 2), np.stack([D] * 2), rig_cam_to_base()[:2])
    assert tr.status == 0 and tr.n_quads > 20
    singles = [po.localize(f, K, D)[0] for f in frames]
    assert tr.n_quads == sum(s.n_quads for s in singles) and tr.n_hyps_generated == sum(s.n_hyps_generated for s in singles)
    assert abs(tr.pos[2] - tb[2]) < 0.1
    # the joint error is a mean over both cameras' projected points: between the best single-camera values' range
    assert np.isfinite(tr.error) and tr.error > 0
