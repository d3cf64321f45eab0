"""Timed CPU leg of bench.py: the oracle's whole-frame restatement of quadDetection()
(src/mantis3.cpp:68-135) run on the host cores over a bounded sample of the bench workload.

TEST / BASELINE INFRASTRUCTURE ONLY.  bench.py's `cpu_baseline` object and `--impl reference`
arm are the only callers; the product path never imports this.
"""
import multiprocessing as mp
import os
import sys
import time

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_HERE)
_FRAMES = {}


def _setup():
    for p in (_ROOT, os.path.join(_ROOT, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)


def _prepare(args):
    """Renders this worker's frames once (untimed) and keeps them in the process."""
    w, h, first, count, seed = args
    _setup()
    import numpy as np

    from frames import K_for, poses
    from oracle import pyoracle as po

    K, D = K_for(w, h), np.zeros(4)
    ps = poses(first + count, seed=seed)[first:]
    _FRAMES[(w, h, first, count, seed)] = ([po.render_grid(R, t, K, D, w, h) for R, t in ps], K, D)
    po.localize(_FRAMES[(w, h, first, count, seed)][0][0], K, D)  # page everything in
    return os.getpid()


def _init(jobs):
    """Pool initializer: every worker renders every job's frames, so that whichever worker a
    job lands on, the timed pass contains only the localization itself."""
    for j in jobs:
        _prepare(j)


def _step(args):
    key, stage, reps = args[:5], args[5], args[6]
    from oracle import pyoracle as po

    frames, K, D = _FRAMES[key]
    ok = 0
    t0 = time.perf_counter()
    for f in frames * reps:
        if stage == "edges":
            po.edge_stage(f, 50)
            ok += 1
        else:
            tr, _m = po.localize(f, K, D)
            ok += tr.status == 0
    return time.perf_counter() - t0, len(frames) * reps, ok


class CpuPath:
    """One single-threaded worker process per core over disjoint frame ranges of the bench
    workload (SURVEY 8d's CPU-baseline plan).  step() = one pass over the sample."""

    def __init__(self, w, h, frames_per_core, cores=None, seed=0x4D414E54):
        from oracle import pyoracle as po

        po.build()
        self.cores = cores or os.cpu_count() or 1
        self.jobs = [(w, h, i * frames_per_core, frames_per_core, seed) for i in range(self.cores)]
        self.pool = mp.get_context("spawn").Pool(self.cores, initializer=_init, initargs=(self.jobs,))
        self.pool.map(_prepare, self.jobs, chunksize=1)  # returns once the workers are up
        self.frames = frames_per_core * self.cores

    def step(self, stage="localize", reps=1):
        """One pass (repeated `reps` times) over the sample.  Returns (wall seconds, frames processed,
        frames with status OK)."""
        t0 = time.perf_counter()
        res = self.pool.map(_step, [j + (stage, reps) for j in self.jobs], chunksize=1)
        wall = time.perf_counter() - t0
        return wall, sum(r[1] for r in res), sum(r[2] for r in res)

    def close(self):
        self.pool.close()
        self.pool.join()
