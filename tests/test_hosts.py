"""The C++ hosts (mantis_b200/bin/mantis3, mantis_server): the reference's two entry points without ROS.
CPU: map parsing as getParameters() does it, loud failure without a device.
GPU: the executables agree with the Python binding on the same frames."""
import os
import re
import subprocess

import numpy as np
import pytest

from frames import K_for, poses
from oracle import pyoracle as po

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "mantis_b200", "bin")


def _build():
    subprocess.check_call(["make", "-C", os.path.join(ROOT, "mantis_b200", "host"), "-s"])


def _write_map_yaml(path):
    w, r, g = po.make_map()
    fmt = lambda pts: ";\n".join("%f, %f, 0" % (p[0], p[1]) for p in pts) + ";"
    with open(path, "w") as f:
        f.write("whiteMap: '%s'\n\nredMap: '%s'\n\ngreenMap: '%s'\n" % (fmt(w), fmt(r), fmt(g)))


def _write_ppm(path, bgr):
    with open(path, "wb") as f:
        f.write(b"P6\n%d %d\n255\n" % (bgr.shape[1], bgr.shape[0]))
        f.write(np.ascontiguousarray(bgr[..., ::-1]).tobytes())


def test_map_yaml_parsing(tmp_path):
    _build()
    y = str(tmp_path / "map.yaml")
    _write_map_yaml(y)
    out = subprocess.check_output([os.path.join(BIN, "mantis3"), "--check-map", y], text=True)
    assert out.startswith("white 646 red 37 green 37 first=(-1.440000 -1.120000 0.000000) last_green=(1.440000 1.440000")


def test_hosts_fail_loudly_without_device(tmp_path):
    import torch

    if torch.cuda.is_available():
        pytest.skip("only meaningful on a box without a GPU")
    _build()
    img = str(tmp_path / "a.ppm")
    _write_ppm(img, np.zeros((8, 8, 3), np.uint8))
    p = subprocess.run([os.path.join(BIN, "mantis3"), "--K", "4,0,4,0,4,4,0,0,1", img], capture_output=True, text=True)
    assert p.returncode == 1 and "no CUDA device" in p.stderr


def _fields(line):
    nums = lambda s: [float(v) for v in s.split()]
    d = {k: v for k, v in re.findall(r"(\w+)=([-\w.+]+)", line)}
    for k, v in re.findall(r"(\w+)=\(([^)]*)\)", line):
        d[k] = nums(v)
    return d


@pytest.mark.gpu
def test_mantis3_matches_binding(ctx, tmp_path):
    W, H = 640, 480
    K, D = K_for(W, H), np.zeros(4)
    frames = [po.render_grid(R, t, K, D, W, H) for R, t in poses(3, seed=4242)]
    paths = []
    for i, f in enumerate(frames):
        paths.append(str(tmp_path / f"f{i}.ppm"))
        _write_ppm(paths[-1], f)
    y = str(tmp_path / "map.yaml")
    _write_map_yaml(y)
    Ks = ",".join(repr(float(v)) for v in K.reshape(-1))
    res, _ = ctx.localize(np.stack(frames), K, D)
    for extra in ([], ["--batch"], ["--map", y]):
        out = subprocess.check_output([os.path.join(BIN, "mantis3"), "--K", Ks, *extra, *paths], text=True)
        lines = [l for l in out.splitlines() if "status=" in l]
        assert len(lines) == len(frames)
        for i, l in enumerate(lines):
            d = _fields(l)
            assert int(d["status"]) == res[i].status and int(d["quads"]) == res[i].n_quads
            assert float(d["error"]) == pytest.approx(res[i].error, rel=1e-8)
            pos = d.get("position", d.get("candidate_position"))
            assert np.allclose(pos, list(res[i].pos), atol=1e-8)
            assert ("pose_estimate" in l) == (res[i].status == 0 and res[i].min_yaw_diff > 4000)


@pytest.mark.gpu
def test_mantis_server_rig_request(ctx, tmp_path):
    """--per-camera: two cameras looking at the same floor, each localized alone; the response is the better
    camera's pose moved into the base frame (world<-base = world<-cam * (base<-cam)^-1)."""
    W, H = 640, 480
    K, D = K_for(W, H), np.zeros(4)
    ps = poses(2, seed=777)
    paths = []
    for i, (R, t) in enumerate(ps):
        paths.append(str(tmp_path / f"c{i}.ppm"))
        _write_ppm(paths[-1], po.render_grid(R, t, K, D, W, H))
    Ks = " ".join(repr(float(v)) for v in K.reshape(-1))
    c2b = [[0.1, 0.0, -0.05, 0, 0, 0, 1], [-0.1, 0.02, -0.05, 0, 0, np.sin(0.2), np.cos(0.2)]]
    req = "request\n"
    for p, cb in zip(paths, c2b):
        req += f"image {p}\nK {Ks}\nD 0 0 0 0\ncam_to_base {' '.join(map(str, cb))}\n"
    req += "delta_pos 0 0 0\ndelta_quat 0 0 0 1\nend\n"
    out = subprocess.run([os.path.join(BIN, "mantis_server"), "--per-camera"], input=req, capture_output=True, text=True,
                         check=True).stdout
    d = _fields(out.strip().splitlines()[-1])
    res, _ = ctx.localize(np.stack([po.render_grid(R, t, K, D, W, H) for R, t in ps]), K, D)
    assert int(d["ok"]) == 1
    assert int(d["num_particles"]) == res[0].n_particles + res[1].n_particles
    ok = [i for i in range(2) if res[i].status == 0 and res[i].min_yaw_diff > 4000]
    assert int(d["valid"]) == (1 if ok else 0)
    if ok:
        best = min(ok, key=lambda i: res[i].error)
        assert int(d["best_camera"]) == best
        # recompute the fusion in numpy
        from frames import rpy_matrix  # noqa: F401  (keeps the helper import explicit)

        def q2m(q):
            x, y, z, w = q
            return np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                             [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                             [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])
        Rwc, twc = q2m(list(res[best].quat_xyzw)), np.array(list(res[best].pos))
        Rbc, tbc = q2m(c2b[best][3:]), np.array(c2b[best][:3])
        Rwb, twb = Rwc @ Rbc.T, twc - Rwc @ Rbc.T @ tbc
        assert np.allclose(d["pose"][:3], twb, atol=1e-7)
        assert np.allclose(q2m(d["pose"][3:]), Rwb, atol=1e-7)
        assert float(d["weight"]) == pytest.approx(res[best].error, rel=1e-8)


@pytest.mark.gpu
def test_mantis_server_joint_rig_request(ctx, tmp_path):
    """A request whose cameras all carry cam_to_base is ONE rig: localized jointly on the device
    (mantis_localize_rig_batch); the response is the base pose and equals the binding's."""
    from frames import rig_cam_to_base, rig_extrinsics

    W, H = 640, 480
    K, D = K_for(W, H), np.zeros(4)
    ext, b2c = rig_extrinsics()[:3], rig_cam_to_base()[:3]
    Rb, tb = poses(1, seed=778)[0]
    frames = np.stack([po.render_grid(Rb @ Rc, tb + Rb @ tc, K, D, W, H) for Rc, tc in ext])
    Ks = " ".join(repr(float(v)) for v in K.reshape(-1))
    req = "request\n"
    for i in range(3):
        p = str(tmp_path / f"r{i}.ppm")
        _write_ppm(p, frames[i])
        req += f"image {p}\nK {Ks}\nD 0 0 0 0\ncam_to_base {' '.join(repr(float(v)) for v in b2c[i])}\n"
    req += "delta_pos 0 0 0\ndelta_quat 0 0 0 1\nend\n"
    out = subprocess.run([os.path.join(BIN, "mantis_server")], input=req, capture_output=True, text=True, check=True).stdout
    d = _fields(out.strip().splitlines()[-1])
    res, _ = ctx.localize_rig(frames[None], np.stack([K] * 3), np.stack([D] * 3), b2c)
    r = res[0]
    assert int(d["ok"]) == 1 and int(d["num_particles"]) == r.n_particles
    assert int(d["valid"]) == int(r.status == 0 and r.min_yaw_diff > 4000)
    if int(d["valid"]):
        assert int(d["best_camera"]) == -1
        assert np.allclose(d["pose"][:3], list(r.pos), atol=1e-8) and np.allclose(d["pose"][3:], list(r.quat_xyzw), atol=1e-8)
        assert float(d["weight"]) == pytest.approx(r.error, rel=1e-8)


@pytest.mark.gpu
def test_device_pool_shards_give_the_single_device_records(tmp_path):
    """mantis3 --devices: one context + one host thread per entry, contiguous shards, records in one array
    (SURVEY 8e).  Two contexts on GPU 0 always; GPUs 0 and 1 when the box has them.  Byte-identical output."""
    import torch

    W, H = 640, 480
    K, D = K_for(W, H), np.zeros(4)
    paths = []
    for i, (R, t) in enumerate(poses(7, seed=4243)):
        paths.append(str(tmp_path / f"p{i}.ppm"))
        _write_ppm(paths[-1], po.render_grid(R, t, K, D, W, H))
    Ks = ",".join(repr(float(v)) for v in K.reshape(-1))
    run = lambda *extra: subprocess.check_output([os.path.join(BIN, "mantis3"), "--K", Ks, *extra, *paths], text=True)
    one = run("--batch")
    assert one.count("status=") == len(paths)
    assert run("--devices", "0,0") == one
    assert run("--devices", "0,0,0") == one
    if torch.cuda.device_count() >= 2:
        assert run("--devices", "0-1") == one
