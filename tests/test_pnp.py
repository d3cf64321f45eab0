"""Gauss-Newton / LM PnP (config 5; the mantis2 generation's computeHypothesis,
include/legacy/mantis2/PoseEstimator.h:16-25,91-118): the oracle against cv2.solvePnP's golden
results (CPU), the CUDA solver against the oracle and against the same golden results (GPU).
Tolerance of the north star: 1e-4 rad rotation, 1 mm translation; measured agreement is ~1e-9."""
import os

import numpy as np
import pytest

from oracle import pyoracle as po

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "pnp_golden.npz"))
ROT_TOL, TRANS_TOL = 1e-4, 1e-3  # the spec
TIGHT = 1e-6                     # what two converged minimisers actually agree to


def rot_angle(r1, r2):
    R1, R2 = po.rodrigues_exp(r1), po.rodrigues_exp(r2)
    return float(np.linalg.norm(R1 - R2) / np.sqrt(2))  # = 2 sin(angle/2) ~ angle


def test_oracle_matches_cv2_on_the_reference_vectors():
    """test/legacy/unit_test.cpp:49-72."""
    o = po.pnp_iterative(G["legacy_obj"], G["legacy_img"], G["legacy_K"])
    assert o["ok"]
    assert rot_angle(o["rvec"], G["legacy_rvec"]) < TIGHT and np.abs(o["tvec"] - G["legacy_tvec"]).max() < TIGHT
    assert rot_angle(o["rvec"], G["legacy_true_rvec"]) < 1e-5 and np.abs(o["tvec"] - G["legacy_true_tvec"]).max() < 1e-5
    assert o["err"] < 1e-3


def test_oracle_matches_cv2_on_200_correspondence_cases():
    for i in range(len(G["obj"])):
        o = po.pnp_iterative(G["obj"][i], G["img"][i], G["K"])
        assert o["ok"]
        assert rot_angle(o["rvec"], G["rvec"][i]) < TIGHT and np.abs(o["tvec"] - G["tvec"][i]).max() < TIGHT, i
        assert 4.0 < o["err"] < 14.0  # sqrt(400 residuals of sigma 0.5 px) ~ 10


def test_oracle_guess_and_second_candidate():
    i = 3
    g = np.concatenate([G["rvec"][i] + 0.05, G["tvec"][i] + 0.05])
    o = po.pnp_iterative(G["obj"][i], G["img"][i], G["K"], guess=g)
    assert rot_angle(o["rvec"], G["rvec"][i]) < TIGHT and np.abs(o["tvec"] - G["tvec"][i]).max() < TIGHT
    o2 = po.pnp_iterative(G["obj"][i], G["img"][i], G["K"], both_planar=True)
    assert o2["err"] <= po.pnp_iterative(G["obj"][i], G["img"][i], G["K"])["err"] + 1e-12
    assert not po.pnp_iterative(G["obj"][i][:3], G["img"][i][:3], G["K"])["ok"]


@pytest.mark.gpu
def test_cuda_pnp_matches_cv2_golden_and_oracle(ctx):
    out = ctx.pnp(G["obj"], G["img"], G["K"])
    for i in range(len(G["obj"])):
        r = out[i]
        assert r.status == 0
        assert rot_angle(r.rvec[:], G["rvec"][i]) < ROT_TOL and np.abs(np.array(r.tvec[:]) - G["tvec"][i]).max() < TRANS_TOL
        o = po.pnp_iterative(G["obj"][i], G["img"][i], G["K"])
        assert rot_angle(r.rvec[:], o["rvec"]) < TIGHT and np.abs(np.array(r.tvec[:]) - o["tvec"]).max() < TIGHT
        assert abs(r.reproj_error - o["err"]) < 1e-6 * o["err"]


@pytest.mark.gpu
def test_cuda_pnp_reference_vectors_guess_flags_and_degenerate(ctx):
    r = ctx.pnp(G["legacy_obj"], G["legacy_img"][None], G["legacy_K"])[0]
    assert r.status == 0 and rot_angle(r.rvec[:], G["legacy_rvec"]) < 1e-5
    assert np.abs(np.array(r.tvec[:]) - G["legacy_tvec"]).max() < 1e-5 and r.reproj_error < 1e-3
    n = 8
    guess = np.concatenate([G["rvec"][:n] + 0.04, G["tvec"][:n] - 0.03], 1)
    out = ctx.pnp(G["obj"][:n], G["img"][:n], G["K"], guess=guess)
    both = ctx.pnp(G["obj"][:n], G["img"][:n], G["K"], both_planar=True)
    for i in range(n):
        assert rot_angle(out[i].rvec[:], G["rvec"][i]) < TIGHT
        assert both[i].reproj_error <= out[i].reproj_error * (1 + 1e-9) and both[i].passes > out[i].passes
    # per-frame K, shared object points, too few points, collinear points
    Ks = np.repeat(G["K"][None], 2, 0)
    out = ctx.pnp(G["obj"][0], G["img"][:2], Ks)
    assert out[0].status == 0 and rot_angle(out[0].rvec[:], G["rvec"][0]) < TIGHT
    assert ctx.pnp(G["obj"][:1, :3], G["img"][:1, :3], G["K"])[0].status == 1
    line = np.zeros((1, 6, 3), np.float32)
    line[0, :, 0] = np.arange(6)
    assert ctx.pnp(line, np.zeros((1, 6, 2), np.float32) + 100, G["K"])[0].status == 1


@pytest.mark.gpu
def test_cuda_pnp_large_batch_is_consistent(ctx):
    """Size-independent property: every copy of a case in a big batch gives the identical answer."""
    reps = 64
    obj = np.tile(G["obj"], (reps, 1, 1))
    img = np.tile(G["img"], (reps, 1, 1))
    out = ctx.pnp(obj, img, G["K"])
    base = np.array([list(out[i].rvec) + list(out[i].tvec) for i in range(len(G["obj"]))])
    allr = np.array([list(o.rvec) + list(o.tvec) for o in out]).reshape(reps, len(G["obj"]), 6)
    assert np.array_equal(allr, np.broadcast_to(base, allr.shape))
