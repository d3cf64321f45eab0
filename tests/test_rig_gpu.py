"""GPU: joint multi-camera localization of a rig (BASELINE.json configs[2]) through mantis_localize_rig_batch against
the oracle's mo_localize_rig (same decisions, same integer scores, base pose within 1e-4 rad / 1 mm), and the
degenerate rig -- one camera at the base origin -- against mantis_localize_batch, byte for byte."""
import numpy as np
import pytest

from frames import D_UNIT, K_for, poses, rig_cam_to_base, rig_extrinsics
from oracle import pyoracle as po
from parity import rot_angle

pytestmark = pytest.mark.gpu


def _render_rig(Rb, tb, K, D, w, h, ext):
    """The cameras of a rig whose BASE sits at (Rb, tb) in the world: camera k at W2B * b2c_k."""
    return np.stack([po.render_grid(Rb @ Rc, tb + Rb @ tc, K, D, w, h) for Rc, tc in ext])


def _compare(res, tr, otr):
    assert res.status == otr.status
    assert (tr.n_contours, tr.n_quads_raw, tr.n_quads) == (otr.n_contours, otr.n_quads_raw, otr.n_quads)
    assert (tr.n_hyps_generated, tr.n_hyps_cluster) == (otr.n_hyps_generated, otr.n_hyps_cluster)
    if otr.status != 0:
        return
    assert tr.first_best_error == otr.first_best.error
    assert tr.optimized_error == otr.optimized.error
    assert list(tr.shift_err) == list(otr.shift_err)
    assert tr.best_rot == otr.best_rot
    assert np.allclose(list(tr.yaw_err), list(otr.yaw_err), rtol=1e-12)
    assert np.abs(np.array(res.pos[:]) - np.array(otr.pos[:])).max() < 1e-3
    assert rot_angle(np.array(res.quat_xyzw[:]), np.array(otr.quat_xyzw[:])) < 1e-4
    assert res.error == pytest.approx(otr.error, rel=1e-12)
    assert res.n_particles == otr.n_scored


def test_rig_of_one_camera_equals_single_camera_path(ctx):
    K, D = K_for(640, 480), np.zeros(4)
    frames = np.stack([po.render_grid(R, t, K, D, 640, 480) for R, t in poses(6, seed=0x4D414E54 + 61)])
    single, _ = ctx.localize(frames, K, D)
    ident = np.array([[0, 0, 0, 0, 0, 0, 1.0]])
    rig, _ = ctx.localize_rig(frames[:, None], K[None], D[None], ident)
    assert bytes(rig) == bytes(single)
    zero = np.zeros((1, 7))   # "not given" = identity
    rig0, _ = ctx.localize_rig(frames[:, None], K[None], D[None], zero)
    assert bytes(rig0) == bytes(single)


@pytest.mark.parametrize("dist", [False, True])
def test_rig_two_cameras_640x480_matches_oracle(ctx, dist):
    K = K_for(640, 480)
    D = D_UNIT if dist else np.zeros(4)
    ext, b2c = rig_extrinsics()[:2], rig_cam_to_base()[:2]
    rigs = np.stack([_render_rig(Rb, tb, K, D, 640, 480, ext) for Rb, tb in poses(3, seed=0x4D414E54 + 62)])
    res, tr = ctx.localize_rig(rigs, np.stack([K] * 2), np.stack([D] * 2), b2c, want_trace=True)
    for i in range(len(rigs)):
        _compare(res[i], tr[i], po.localize_rig(rigs[i], np.stack([K] * 2), np.stack([D] * 2), b2c))


def test_config2_four_camera_rig_1280x720_matches_oracle(ctx):
    """BASELINE.json configs[2] / SURVEY 8d config 3: 4 cameras (down, 30 degrees fore, aft, side), 1280x720."""
    W, H = 1280, 720
    K, D = K_for(W, H), np.zeros(4)
    ext, b2c = rig_extrinsics(), rig_cam_to_base()
    ps = poses(3, seed=0x4D414E54 + 3)
    rigs = np.stack([_render_rig(Rb, tb, K, D, W, H, ext) for Rb, tb in ps])
    Ks, Ds = np.stack([K] * 4), np.stack([D] * 4)
    res, tr = ctx.localize_rig(rigs, Ks, Ds, b2c, want_trace=True)
    for i in range(len(rigs)):
        otr = po.localize_rig(rigs[i], Ks, Ds, b2c)
        assert otr.status == 0
        _compare(res[i], tr[i], otr)
        # and the base comes back where it was put (modulo the grid's symmetries the yaw step resolves by colour)
        assert abs(res[i].pos[2] - ps[i][1][2]) < 0.1


def test_rig_batches_are_chunked_on_rig_boundaries(ctx, monkeypatch):
    """More rigs than a chunk holds: the records must not depend on the chunking."""
    K, D = K_for(640, 480), np.zeros(4)
    ext, b2c = rig_extrinsics()[:3], rig_cam_to_base()[:3]
    rigs = np.stack([_render_rig(Rb, tb, K, D, 640, 480, ext) for Rb, tb in poses(5, seed=0x4D414E54 + 63)])
    Ks, Ds = np.stack([K] * 3), np.stack([D] * 3)
    whole, _ = ctx.localize_rig(rigs, Ks, Ds, b2c)
    parts = b"".join(bytes(ctx.localize_rig(rigs[i:i + 1], Ks, Ds, b2c)[0]) for i in range(len(rigs)))
    assert bytes(whole) == parts
    assert all(r.status == 0 for r in whole)
