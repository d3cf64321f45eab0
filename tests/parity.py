"""Whole-batch parity of the CUDA path against the oracle, frame by frame (TEST INFRASTRUCTURE).

`run()` pushes n frames through mantis_edges_batch / mantis_quads_batch / mantis_localize_batch and,
in parallel on the host cores, through the oracle (one spawned worker per core reading the frames from a
shared-memory file), then records per frame: Canny / morphology map equality, quad-list equality, the
equality of every discrete decision and integer score of the trace, and |dpos|, drot against the oracle's
pose.  SURVEY.md 8(d): every frame bit-exact edges + equal quad sets; poses within 1e-4 rad / 1 mm.
Used by tests/test_parity_full_gpu.py and by bench.py's `parity_sample` key (as the checker, outside the
timed region).
"""
import multiprocessing as mp
import os
import sys
import tempfile

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_HERE)
POS_TOL = 1e-3    # 1 mm   (BASELINE.json north_star)
ROT_TOL = 1e-4    # 1e-4 rad

_W = {}


def _setup():
    for p in (_ROOT, _HERE):
        if p not in sys.path:
            sys.path.insert(0, p)


def _worker_init(path, shape, K, D, edge_paths):
    _setup()
    _W["frames"] = np.load(path, mmap_mode="r")
    assert _W["frames"].shape == tuple(shape)
    _W["K"], _W["D"] = K, D
    _W["canny"] = np.load(edge_paths[0], mmap_mode="r") if edge_paths else None
    _W["morph"] = np.load(edge_paths[1], mmap_mode="r") if edge_paths else None


def _oracle_frame(i):
    from oracle import pyoracle as po

    f = np.ascontiguousarray(_W["frames"][i])
    tr, _ = po.localize(f, _W["K"], _W["D"])
    rec = {
        "i": i, "status": tr.status, "n_contours": tr.n_contours, "n_quads_raw": tr.n_quads_raw, "n_quads": tr.n_quads,
        "n_hyps_generated": tr.n_hyps_generated, "n_hyps_cluster": tr.n_hyps_cluster,
        "corners": np.array([list(tr.quad_corners[q]) for q in range(tr.n_quads)], np.int32).reshape(-1, 8),
        "test_px": np.array([list(tr.quad_test_px[q]) for q in range(tr.n_quads)], np.float64).reshape(-1, 8),
        "first_best": tr.first_best.error, "optimized": tr.optimized.error, "shift_err": list(tr.shift_err),
        "yaw_err": list(tr.yaw_err), "best_rot": tr.best_rot, "min_yaw_diff": tr.min_yaw_diff,
        "pos": list(tr.pos), "quat": list(tr.quat_xyzw), "error": tr.error, "n_scored": tr.n_scored,
    }
    if _W["canny"] is not None:
        oc, om = po.edge_stage(f, 50)
        rec["canny_equal"] = bool(np.array_equal(oc, _W["canny"][i]))
        rec["morph_equal"] = bool(np.array_equal(om, _W["morph"][i]))
    return rec


def rot_angle(q1, q2):
    """Angle of the rotation between two unit quaternions (x y z w), accurate near 0: 2 atan2(|v|, |w|) of q1^-1 q2."""
    x1, y1, z1, w1 = q1
    x2, y2, z2, w2 = q2
    w = w1 * w2 + x1 * x2 + y1 * y2 + z1 * z2
    v = np.array([w1 * x2 - x1 * w2 - y1 * z2 + z1 * y2, w1 * y2 + x1 * z2 - y1 * w2 - z1 * x2,
                  w1 * z2 - x1 * y2 + y1 * x2 - z1 * w2])
    return 2 * float(np.arctan2(np.linalg.norm(v), abs(w)))


def compare_frame(res, tr, fq, o):
    """One frame: GPU (result, trace, quads record) against the oracle record `o`.  Returns a dict of findings."""
    out = {"i": o["i"], "status": (res.status, o["status"])}
    out["counts_equal"] = (res.status == o["status"]
                           and (tr.n_contours, tr.n_quads_raw, tr.n_quads) == (o["n_contours"], o["n_quads_raw"], o["n_quads"])
                           and (tr.n_hyps_generated, tr.n_hyps_cluster) == (o["n_hyps_generated"], o["n_hyps_cluster"]))
    if fq is not None:
        nq = fq.n_quads
        gc = np.array([list(fq.corners[q]) for q in range(nq)], np.int32).reshape(-1, 8)
        gt = np.array([list(fq.test_px[q]) for q in range(nq)], np.float64).reshape(-1, 8)
        out["quads_equal"] = bool(nq == o["n_quads"] and np.array_equal(gc, o["corners"]) and np.array_equal(gt, o["test_px"]))
    for k in ("canny_equal", "morph_equal"):
        if k in o:
            out[k] = o[k]
    if o["status"] != 0 or res.status != 0:
        out["trace_equal"] = out["counts_equal"]
        out["dpos"], out["drot"] = 0.0, 0.0
        return out
    out["trace_equal"] = bool(out["counts_equal"] and tr.first_best_error == o["first_best"]
                              and tr.optimized_error == o["optimized"] and list(tr.shift_err) == o["shift_err"]
                              and tr.best_rot == o["best_rot"]
                              and np.allclose(list(tr.yaw_err), o["yaw_err"], rtol=1e-12)
                              and res.n_particles == o["n_scored"])
    out["dpos"] = float(np.abs(np.array(res.pos[:]) - np.array(o["pos"])).max())
    out["drot"] = float(rot_angle(np.array(res.quat_xyzw[:]), np.array(o["quat"])))
    out["derr"] = float(abs(res.error - o["error"]) / max(abs(o["error"]), 1e-300))
    if not out["trace_equal"]:  # where the two paths part
        if tr.first_best_error != o["first_best"]:
            out["first_difference"] = f"first_best {tr.first_best_error!r} vs {o['first_best']!r}"
        elif tr.optimized_error != o["optimized"]:
            out["first_difference"] = f"optimized {tr.optimized_error!r} vs {o['optimized']!r}"
        elif list(tr.shift_err) != o["shift_err"]:
            out["first_difference"] = "shift_err"
        elif tr.best_rot != o["best_rot"]:
            out["first_difference"] = f"best_rot {tr.best_rot} vs {o['best_rot']}"
        else:
            out["first_difference"] = "yaw_err / n_scored"
    return out


def summarize(findings):
    n = len(findings)
    s = {
        "frames": n,
        "status_ok": sum(1 for f in findings if f["status"] == (0, 0)),
        "counts_equal": sum(f["counts_equal"] for f in findings),
        "identical": sum(f["trace_equal"] for f in findings),
        "within_tol": sum(f["counts_equal"] and f["dpos"] < POS_TOL and f["drot"] < ROT_TOL for f in findings),
        "max_dpos": max((f["dpos"] for f in findings), default=0.0),
        "max_drot": max((f["drot"] for f in findings), default=0.0),
    }
    for k in ("quads_equal", "canny_equal", "morph_equal"):
        if findings and k in findings[0]:
            s[k] = sum(f[k] for f in findings)
    s["offenders"] = [
        {k: v for k, v in f.items() if k in ("i", "status", "dpos", "drot", "first_difference", "counts_equal")}
        for f in findings if not (f["counts_equal"] and f["dpos"] < POS_TOL and f["drot"] < ROT_TOL)][:32]
    return s


def run(ctx, frames_host, K, D, want_edges=True, want_quads=True, procs=None, chunk=512):
    """frames_host: uint8 [n, H, W, 3] on the host (the bytes both sides see).  Returns (summary, findings)."""
    _setup()
    from oracle import pyoracle as po

    po.build()
    n, H, W, _ = frames_host.shape
    K = np.ascontiguousarray(K, np.float32)
    D = np.ascontiguousarray(D, np.float64)
    shm = "/dev/shm" if os.path.isdir("/dev/shm") else None
    tmp = tempfile.mkdtemp(prefix="mantis_parity_", dir=shm)
    fpath = os.path.join(tmp, "frames.npy")
    np.save(fpath, frames_host)
    edge_paths = None
    try:
        # ---- CUDA path ----
        res_all, tr_all, fq_all = [], [], []
        canny = morph = None
        if want_edges:
            edge_paths = (os.path.join(tmp, "canny.npy"), os.path.join(tmp, "morph.npy"))
            canny = np.lib.format.open_memmap(edge_paths[0], mode="w+", dtype=np.uint8, shape=(n, H, W))
            morph = np.lib.format.open_memmap(edge_paths[1], mode="w+", dtype=np.uint8, shape=(n, H, W))
        for s in range(0, n, chunk):
            part = frames_host[s:s + chunk]
            res, tr = ctx.localize(part, K, D, want_trace=True)
            res_all.extend(res)
            tr_all.extend(tr)
            if want_quads:
                fq_all.extend(ctx.quads(part))
            if want_edges:
                c, m = ctx.edges(part)
                canny[s:s + len(part)] = c
                morph[s:s + len(part)] = m
        if want_edges:
            canny.flush()
            morph.flush()
            del canny, morph
        # ---- oracle, one worker per core ----
        procs = procs or os.cpu_count() or 1
        with mp.get_context("spawn").Pool(procs, initializer=_worker_init,
                                          initargs=(fpath, frames_host.shape, K, D, edge_paths)) as pool:
            orecs = pool.map(_oracle_frame, range(n), chunksize=max(1, n // (procs * 8)))
        findings = [compare_frame(res_all[i], tr_all[i], fq_all[i] if want_quads else None, orecs[i]) for i in range(n)]
        return summarize(findings), findings
    finally:
        for f in os.listdir(tmp):
            os.unlink(os.path.join(tmp, f))
        os.rmdir(tmp)
