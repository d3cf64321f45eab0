#!/usr/bin/env python
"""bench.py -- localized frames/s of the mantis per-frame hot path on B200 (BASELINE.json's metric).

A step = one pass of mantis_localize_batch() (everything quadDetection() does, src/mantis3.cpp:68-135)
over one batch of synthetic rendered grid frames (BASELINE.json configs[1]: 640x480, batch 4096 per GPU).

  value     frames/s with the frames already resident in HBM (device pointers through the C ABI),
            CUDA events on the library's stream, max over ranks
  e2e       the same call with HOST (pinned) frames: H2D of every frame and D2H of every result inside
            the timed region
  roofline  image stage (k_edge_candidates + k_edge_link: BGR in -> final edge map, 4 B/px algorithmic),
            timed live with CUDA events inside the timed steps, against MEASURED_PEAKS.json's HBM figure
  cpu_baseline  the CPU oracle (a C restatement of the reference's path, kind "port") on the box's host
            cores over a bounded sample of the same frames (rank 0, N=1 only)
  per_rank  ms per step of every rank (min / median / max): what the max-over-ranks `value` hides
  parity_sample, hd720_frames_per_s, rig_4x720p_rigs_per_s, hd1080_noise_frames_per_s, d_nonzero_frames_per_s,
  pnp_frames_per_s   (N = 1) the other BASELINE.json configs and a checked sample, tools/bench_extras.py

  --impl reference   times that CPU path alone (the reference has no GPU code and cannot be built here:
                     it needs ROS + OpenCV C++).

Multi-GPU: one process per GPU (torchrun), frames sharded by rank, no data-path collective
(torch.distributed is used only for the barrier and the max-over-ranks of the timings).
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "tools"))

METRIC = "localized frames/sec at 640x480"
UNIT = "frames/s"
W, H = 640, 480
SEED = 0x4D414E54 + 2  # "MANT" + config id


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=4096, help="frames per GPU per step")
    ap.add_argument("--cpu-frames-per-core", type=int, default=4)
    ap.add_argument("--cpu-reps", type=int, default=40, help="passes over the CPU sample inside its timer")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the other configs' keys and the parity sample (N = 1 only)")
    ap.add_argument("--parity-frames", type=int, default=512, help="frames of the workload checked against the oracle")
    return ap.parse_args()


def workload(batch, n_gpus):
    return {"workload": f"synthetic rendered grid frames {W}x{H}, batch {batch} per GPU (BASELINE.json configs[1])",
            "width": W, "height": H, "batch_per_gpu": batch, "global_batch": batch * n_gpus,
            "camera": "fx=fy=320 cx=320 cy=240, D=0 (fisheye model)", "parallelism": f"frame-sharded x{n_gpus}, no collective",
            "l2_policy": f"inputs larger than L2 ({batch * W * H * 3 / 1e6:.0f} MB of frames per GPU per step)"}


# ---------------------------------------------------------------------------------------------
class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md's clocks line)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.index)], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def mark(self):
        """Call at the start of the timed region: samples before it (the warm-up steps, same load) are only used
        when nvidia-smi delivered none inside the region (a short region, or eight nvidia-smi processes at once)."""
        self.first = len(self.rows)

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        first = getattr(self, "first", 0)
        window = "timed region"
        rows = self.rows[first:]
        if not rows:
            rows, window = self.rows[-5:], "warm-up steps (no sample arrived inside the timed region)"
        for r in rows:
            if len(r) < 9:
                continue
            try:
                sm.append(float(r[1]))
                mx.append(float(r[2]))
            except ValueError:
                continue
            for k, name in enumerate(names):
                if r[5 + k].lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm), "window": window}


def hbm_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except (OSError, KeyError, ValueError):
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


# ---------------------------------------------------------------------------------------------
def run_reference(args):
    """The reference's own CPU implementation of the path, on all host cores.  The reference cannot be
    compiled here (ROS + OpenCV C++), so this is the oracle port (kind "port")."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle.cpu_bench import CpuPath

    fpc = max(1, args.cpu_frames_per_core // 2)
    cp = CpuPath(W, H, fpc, seed=SEED)
    for _ in range(args.warmup):
        cp.step()
    t0 = time.perf_counter()
    frames = ok = 0
    for _ in range(args.steps):
        _s, n, k = cp.step()
        frames += n
        ok += k
    dt = time.perf_counter() - t0
    cp.close()
    v = frames / dt
    sample = f"{cp.frames} frames per step ({fpc} per core), the first frames of the bench workload's pose stream"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8 image stages, int32 contours, f64 pose/scoring", "data": "synthetic",
        "config": workload(args.batch, args.gpus),
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cp.cores, "kind": "port", "sample": sample,
                         "frames_ok": ok, "frames": frames},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}))


# ---------------------------------------------------------------------------------------------
def main():
    args = parse()
    if args.impl == "reference":
        return run_reference(args)

    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))

    # CPU baseline first (rank 0, N=1): spawned workers, before this process touches CUDA
    cpu = None
    if world == 1 and args.gpus == 1 and not args.no_cpu_baseline:
        from oracle.cpu_bench import CpuPath

        cp = CpuPath(W, H, args.cpu_frames_per_core, seed=SEED)
        cp.step()
        secs, n, ok = cp.step(reps=args.cpu_reps)
        es, en, _ = cp.step("edges", reps=args.cpu_reps)
        cp.close()
        cpu = {"value": n / secs, "unit": UNIT, "cores": cp.cores, "kind": "port",
               "sample": f"{n} frame localizations ({args.cpu_frames_per_core} distinct frames per core x {args.cpu_reps} passes, "
                         f"one single-threaded process per core), "
                         f"the first frames of the bench workload's pose stream; {secs:.1f} s",
               "frames_ok": ok, "image_stage_frames_per_s": en / es}

    import numpy as np
    import torch

    import mantis_b200
    from frames import K_for, poses

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: libmantis_cuda has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        # several ranks share the host: keep each one (its pinned staging memory, its copy-issuing thread) on the
        # CPUs next to its GPU, or the end-to-end arm crosses the socket interconnect
        try:
            import pynvml

            pynvml.nvmlInit()
            pynvml.nvmlDeviceSetCpuAffinity(pynvml.nvmlDeviceGetHandleByIndex(local))
        except Exception:  # noqa: BLE001 -- affinity is an optimisation, never a requirement
            pass
    dist = None
    if world > 1:
        import torch.distributed as dist

        # the contract is ONE JSON line on stdout: NCCL prints its version banner there when NCCL_DEBUG=VERSION
        if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def all_ranks(x):
        """x of every rank (rank order)."""
        if dist is None:
            return [x]
        t = [torch.zeros(1, dtype=torch.float64, device="cuda") for _ in range(world)]
        dist.all_gather(t, torch.tensor([x], dtype=torch.float64, device="cuda"))
        return [float(v.item()) for v in t]

    B = args.batch
    K, D = K_for(W, H), np.zeros(4)
    ctx = mantis_b200.Context(local)
    dev = torch.device("cuda", local)
    buf = torch.empty((B, H, W, 3), dtype=torch.uint8, device=dev)
    # rank r renders frames [r*B, (r+1)*B) of one pose stream; rank 0's first frames are the CPU sample
    ps = poses((rank + 1) * B, seed=SEED)[rank * B:]
    ctx.render_grid_device(ps, W, H, K, D, buf.data_ptr())
    torch.cuda.synchronize()
    dframes = ctx.make_frames([buf.data_ptr() + i * H * W * 3 for i in range(B)], W, H, K, D)
    stream = torch.cuda.ExternalStream(ctx.stream, device=dev)

    # ---- device-resident arm ----
    sampler = ClockSampler(local)
    sampler.start()  # before the warm-up: nvidia-smi needs a moment to deliver its first line
    for _ in range(args.warmup):
        res, _ = ctx.localize_frames(dframes, mantis_b200.MEM_DEVICE)
    status = np.bincount([r.status for r in res], minlength=5).tolist()
    barrier()
    sampler.mark()
    l0 = ctx.launch_count
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        ctx.localize_frames(dframes, mantis_b200.MEM_DEVICE)
    e1.record(stream)
    barrier()
    launches = ctx.launch_count - l0
    clocks = sampler.stop()
    ms_rank = e0.elapsed_time(e1)
    ms_all = all_ranks(ms_rank)
    ms = max(ms_all)
    value = world * B * args.steps / (ms * 1e-3)
    per_rank = {"ms_per_step": [m / args.steps for m in ms_all], "min": min(ms_all) / args.steps,
                "median": statistics.median(ms_all) / args.steps, "max": max(ms_all) / args.steps}

    # ---- per-stage device times: the same K steps once more with the library's stage events on (CUDA events
    # on the library's stream between the stages).  The library runs one lane by default, so this pass and
    # the timed pass above execute the same launches; with MANTIS_LANES >= 2 the timed pass would overlap
    # lanes and this one would not.  The image-stage kernels are timed here.
    ctx.set_profiling(True)
    ctx.localize_frames(dframes, mantis_b200.MEM_DEVICE)  # untimed: the single-lane path sizes its own scratch here
    ctx.stage_ms()
    p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    p0.record(stream)
    for _ in range(args.steps):
        ctx.localize_frames(dframes, mantis_b200.MEM_DEVICE)
    p1.record(stream)
    torch.cuda.synchronize()
    ms_single = p0.elapsed_time(p1)
    stage = ctx.stage_ms()
    ctx.set_profiling(False)

    # ---- roofline of the image stage, from the events recorded inside the timed steps ----
    peak, peak_src = hbm_peak()
    img_ms = (stage["edge_candidates"] + stage["edge_link"]) / args.steps
    alg_bytes = 4.0 * W * H * B  # read 3 B/px BGR + write 1 B/px edge map, per frame
    achieved = alg_bytes / (img_ms * 1e-3) / 1e9
    roofline = {"bound": "hbm", "kernel": "image stage = k_edge_candidates + k_edge_link (2 launches per chunk of up to 4096 frames)",
                "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
                "peak_source": peak_src, "algorithmic_bytes_per_frame": 4 * W * H,
                "image_stage_frames_per_s": B / (img_ms * 1e-3),
                "share_of_step": img_ms / (ms_single / args.steps),
                "timing": "CUDA events around K1 and K2 on the library stream, in a second pass of the same K steps "
                          "(same launches as the headline pass, plus the stage events)"}
    prof = os.path.join(ROOT, "profiles", "image_stage_traffic.json")
    if os.path.exists(prof):
        try:
            with open(prof) as f:
                per_frame = json.load(f).get("dram_bytes_per_frame")
                roofline["traffic_bytes_per_frame"] = per_frame  # ncu dram read + write of K1 + K2, per frame
                roofline["traffic"] = per_frame * B if per_frame else None  # per launch pair, like `achieved`'s numerator
                roofline["algorithmic_bytes_per_launch"] = alg_bytes
        except (OSError, ValueError):
            pass

    # ---- end to end: host (pinned) frames through the same C-ABI call ----
    e2e = None
    host = None
    if not args.no_e2e:
        nb = B
        try:
            host = torch.empty((nb, H, W, 3), dtype=torch.uint8, pin_memory=True)
        except RuntimeError:
            nb = min(B, 1024)
            host = torch.empty((nb, H, W, 3), dtype=torch.uint8, pin_memory=True)
        host.copy_(buf[:nb])
        torch.cuda.synchronize()
        hframes = ctx.make_frames([host.data_ptr() + i * H * W * 3 for i in range(nb)], W, H, K, D)
        for _ in range(min(args.warmup, 2)):
            ctx.localize_frames(hframes, mantis_b200.MEM_HOST)
        ksteps = max(1, args.steps)
        barrier()
        t0 = time.perf_counter()
        for _ in range(ksteps):
            hres, _ = ctx.localize_frames(hframes, mantis_b200.MEM_HOST)
        torch.cuda.synchronize()
        dt_all = all_ranks(time.perf_counter() - t0)
        dt = max(dt_all)
        e2e = {"value": world * nb * ksteps / dt, "unit": UNIT, "h2d_bytes_per_step": nb * H * W * 3,
               "d2h_bytes_per_step": nb * 72 + nb * 16, "frames_per_gpu_per_step": nb, "steps": ksteps,
               "h2d_gb_per_s_aggregate": world * nb * H * W * 3 * ksteps / dt / 1e9,
               "per_rank_s_per_step": [d / ksteps for d in dt_all],
               "timer": "host wall clock around the synchronous C-ABI call, max over ranks"}
        assert [r.status for r in hres[:64]] == [r.status for r in res[:64]]

    # ---- the other configs and the parity sample (N = 1; each timed with CUDA events in this same run) ----
    extras = {}
    if world == 1 and not args.no_extras:
        import bench_extras as bx
        from frames import D_UNIT

        host = None   # (the pinned staging buffer of the e2e arm is no longer needed)
        torch.cuda.empty_cache()
        if args.parity_frames > 0:
            extras["parity_sample"] = bx.parity_sample(ctx, buf, W, H, K, D, min(args.parity_frames, B))
        del dframes
        buf = None
        torch.cuda.empty_cache()
        extras["d_nonzero_frames_per_s"] = bx.localize_rate(ctx, W, H, B, D_UNIT, SEED + 100)
        extras["d_nonzero_frames_per_s"]["camera"] = "fx=fy=320 cx=320 cy=240, D = test/unit_test.cpp:144"
        extras["hd720_frames_per_s"] = bx.localize_rate(ctx, 1280, 720, 1024, np.zeros(4), SEED + 101)
        extras["rig_4x720p_rigs_per_s"] = bx.rig_rate(ctx, 1280, 720, 256, SEED + 102)
        extras["hd1080_noise_frames_per_s"] = bx.localize_rate(ctx, 1920, 1080, 512, np.zeros(4), SEED + 103, degrade=True)
        extras["pnp_frames_per_s"] = bx.pnp_rate(ctx)

    if rank == 0:
        out = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
               "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
               "vs_baseline": None, "dtype": "u8 image stages, int32 contours, f64 pose/scoring", "data": "synthetic",
               "config": workload(B, world), "clocks": clocks, "e2e": e2e, "gpu_launches": launches,
               "roofline": roofline, "per_rank": per_rank,
               "stage_ms_per_step": {k: v / args.steps for k, v in stage.items()},
               "single_lane_ms_per_step": ms_single / args.steps, "lanes": int(os.environ.get("MANTIS_LANES", "1")),
               "frame_status_histogram": dict(zip(["ok", "no_quads", "no_hypotheses", "yaw_failed", "overflow"], status))}
        if cpu is not None:
            out["cpu_baseline"] = cpu
        out.update(extras)
        print(json.dumps(out))
    ctx.close()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
