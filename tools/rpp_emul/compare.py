"""Diagnostic: device RPP (host build, tools/rpp_emul/emul.cpp) against the oracle's mo_rpp on the quads of
seeded rendered frames: per (quad, ordering) the iteration counts of every ObjPose run and the final pose.
  python tools/rpp_emul/compare.py [n_frames] [seed]
"""
import ctypes
import os
import subprocess
import sys
from multiprocessing import Pool

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from frames import K_for, poses  # noqa: E402
from oracle import pyoracle as po  # noqa: E402


def build():
    so = os.path.join(HERE, "librpp_emul.so")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-mfma", *os.environ.get("EMUL_FLAGS", "").split(), "-shared", "-fPIC", "-o", so,
                           os.path.join(HERE, "emul.cpp")])
    return ctypes.CDLL(so)


def quads_of(args):
    i, (R, t), W, H = args
    K, D = K_for(W, H), np.zeros(4)
    tr, _ = po.localize(po.render_grid(R, t, K, D, W, H), K, D)
    return i, np.array([list(tr.quad_test_norm[q]) for q in range(tr.n_quads)])


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 64
    seed = int(sys.argv[2], 0) if len(sys.argv) > 2 else 0x4D414E54 + 2
    W, H = 640, 480
    emul = build()
    dp = lambda a: a.ctypes.data_as(ctypes.c_void_p)
    olib = po.lib()
    dbg_it = (ctypes.c_int * 8).in_dll(olib, "mo_rpp_debug_iters")
    dbg_n = ctypes.c_int.in_dll(olib, "mo_rpp_debug_n")
    with Pool(os.cpu_count()) as pool:
        allq = pool.map(quads_of, [(i, p, W, H) for i, p in enumerate(poses(n, seed))])
    s = 0.16
    poss = [[(s, s), (-s, s), (-s, -s), (s, -s)], [(s, -s), (-s, -s), (-s, s), (s, s)]]
    n_runs = n_diff_it = n_quads = n_bad = 0
    frames_bad = set()
    for fi, qs in allq:
        for qi, q in enumerate(qs):
            for e in range(2):
                model = np.zeros((3, 4))
                ip = np.ones((3, 4))
                for k in range(4):
                    model[0, k], model[1, k] = poss[e][k]
                    ip[0, k], ip[1, k] = q[2 * k], q[2 * k + 1]
                ok, Ro, to, it, oe, ie = po.rpp(model, ip)
                o_its = [dbg_it[k] for k in range(dbg_n.value)]
                R9, t3, its = np.zeros(9), np.zeros(3), (ctypes.c_int * 8)()
                nit, eo, ei = ctypes.c_int(), ctypes.c_double(), ctypes.c_double()
                rays = np.ascontiguousarray(q, np.float64)
                ok2 = emul.emul_rpp(dp(rays), ctypes.c_double(s), e, dp(R9), dp(t3), its, ctypes.byref(nit), ctypes.byref(eo),
                                    ctypes.byref(ei))
                d_its = [its[k] for k in range(nit.value)]
                n_quads += 1
                n_runs += len(o_its)
                # the oracle counts AbsKernel calls, the device counts stop tests
                if bool(ok2) != ok:
                    print(f"frame {fi} quad {qi} ord {e}: ok {ok} vs {bool(ok2)}")
                    n_bad += 1
                    frames_bad.add(fi)
                    continue
                if not ok:
                    continue
                dR = np.abs(R9.reshape(3, 3) - Ro).max()
                dt = np.abs(t3 - to).max()
                if dR > 1e-9 or dt > 1e-9:
                    n_bad += 1
                    frames_bad.add(fi)
                    print(f"frame {fi} quad {qi} ord {e}: dR {dR:.2e} dt {dt:.2e} oracle its {o_its} device its {d_its} "
                          f"obj_err {oe:.3e}/{eo.value:.3e}")
    print(f"{n} frames, {n_quads} (quad, ordering) solves, {n_bad} differ beyond 1e-9 in {len(frames_bad)} frames")


if __name__ == "__main__":
    main()
