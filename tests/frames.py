"""Seeded synthetic grid frames + camera fixtures shared by tests, golden generators and bench."""
import numpy as np

SEED = 0x4D414E54

# test/unit_test.cpp:144-145 of the reference (fisheye D, K of its 1280x1024 camera)
K_UNIT = np.array([[323.1511535644531, 0.0, 642.658203125], [0.0, 322.78955078125, 501.5538330078125],
                   [0.0, 0.0, 1.0]], np.float32)
D_UNIT = np.array([0.0029509200248867273, -0.009944040328264236, 0.005587350111454725, -0.00205406011082232])
# test/grid1.jpg (907x505) carries no intrinsics in the reference; fixture choice: the unit-test
# camera scaled by 907/1280 and centred, with its distortion coefficients
K_GRID1 = np.array([[229.0, 0.0, 453.5], [0.0, 228.75, 252.5], [0.0, 0.0, 1.0]], np.float32)


def K_for(w, h):
    """fx = fy = w/2, principal point at the centre, D = 0 (equidistant fisheye)."""
    return np.array([[w / 2.0, 0, w / 2.0], [0, w / 2.0, h / 2.0], [0, 0, 1]], np.float32)


def rpy_matrix(roll, pitch, yaw):
    cr, sr, cp, sp, cy, sy = np.cos(roll), np.sin(roll), np.cos(pitch), np.sin(pitch), np.cos(yaw), np.sin(yaw)
    Rz = np.array([[cy, -sy, 0], [sy, cy, 0], [0, 0, 1]])
    Ry = np.array([[cp, 0, sp], [0, 1, 0], [-sp, 0, cp]])
    Rx = np.array([[1, 0, 0], [0, cr, -sr], [0, sr, cr]])
    return Rz @ Ry @ Rx


def random_pose(rng):
    """Camera pose in the world (w2c): height U(0.8,2), roll/pitch N(0,0.1), yaw U(0,2pi), xy U(-1,1)."""
    R = rpy_matrix(0, 0, rng.uniform(0, 2 * np.pi)) @ np.diag([1.0, -1.0, -1.0]) @ rpy_matrix(
        rng.normal(0, 0.1), rng.normal(0, 0.1), 0)
    t = np.array([rng.uniform(-1, 1), rng.uniform(-1, 1), rng.uniform(0.8, 2.0)])
    return R, t


def poses(n, seed=SEED):
    rng = np.random.default_rng(seed)
    return [random_pose(rng) for _ in range(n)]


def quat_xyzw(R):
    """Rotation matrix -> unit quaternion (x, y, z, w), w >= 0."""
    w = np.sqrt(max(0.0, 1 + R[0, 0] + R[1, 1] + R[2, 2])) / 2
    if w > 1e-6:
        q = np.array([(R[2, 1] - R[1, 2]) / (4 * w), (R[0, 2] - R[2, 0]) / (4 * w), (R[1, 0] - R[0, 1]) / (4 * w), w])
    else:
        i = int(np.argmax(np.diag(R)))
        j, k = (i + 1) % 3, (i + 2) % 3
        s = np.sqrt(max(0.0, 1 + R[i, i] - R[j, j] - R[k, k])) * 2
        q = np.zeros(4)
        q[i] = s / 4
        q[j] = (R[j, i] + R[i, j]) / s
        q[k] = (R[k, i] + R[i, k]) / s
        q[3] = (R[k, j] - R[j, k]) / s
    return q / np.linalg.norm(q)


def rig_extrinsics():
    """four cameras on the base -- straight down, 30 degrees fore, 30 degrees aft, 30 degrees to the
    side -- each 5 cm off the base origin.  Returns [(R_b2c, t_b2c)] = the pose of each camera in the base frame (the
    base frame is itself a down-looking camera frame: x right, y back, z down)."""
    a = np.deg2rad(30)
    out = [(np.eye(3), np.array([0.0, 0.0, 0.0])),
           (rpy_matrix(a, 0, 0), np.array([0.0, -0.05, 0.0])),
           (rpy_matrix(-a, 0, 0), np.array([0.0, 0.05, 0.0])),
           (rpy_matrix(0, a, 0), np.array([0.05, 0.0, 0.0]))]
    return out


def rig_cam_to_base():
    return np.array([np.concatenate([t, quat_xyzw(R)]) for R, t in rig_extrinsics()])
