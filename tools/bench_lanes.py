"""Device-resident localization throughput for the current MANTIS_LANES setting (wall clock around the C-ABI call)."""
import sys, time, os
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import mantis_b200
from frames import K_for, poses
W, H, N = 640, 480, int(sys.argv[1]) if len(sys.argv) > 1 else 4096
K, D = K_for(W, H), np.zeros(4)
ctx = mantis_b200.Context(0)
buf = torch.empty((N, H, W, 3), dtype=torch.uint8, device="cuda:0")
ctx.render_grid_device(poses(N, seed=1234), W, H, K, D, buf.data_ptr())
fr = ctx.make_frames([buf.data_ptr() + i * H * W * 3 for i in range(N)], W, H, K, D)
torch.cuda.synchronize()
for _ in range(2):
    ctx.localize_frames(fr, mantis_b200.MEM_DEVICE)
ts = []
for _ in range(5):
    t0 = time.perf_counter(); ctx.localize_frames(fr, mantis_b200.MEM_DEVICE); ts.append(time.perf_counter() - t0)
print("MANTIS_LANES=%s N=%d: best %.1f ms -> %.0f frames/s" % (os.environ.get("MANTIS_LANES", "default"), N, min(ts) * 1e3, N / min(ts)))
