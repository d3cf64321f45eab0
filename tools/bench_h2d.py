"""Pinned host -> device copy bandwidth of the box: the ceiling of bench.py's end-to-end number.
Alone: one GPU.  Under torchrun (one rank per GPU) every rank copies AT THE SAME TIME -- same barrier + max-over-ranks
timing as bench.py -- and rank 0 prints the aggregate, i.e. what the host (memory + PCIe fabric) can feed N GPUs.
  python tools/bench_h2d.py
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29517 tools/bench_h2d.py"""
import json
import os
import time

import torch

rank, local, world = (int(os.environ.get(k, d)) for k, d in (("RANK", "0"), ("LOCAL_RANK", "0"), ("WORLD_SIZE", "1")))
torch.cuda.set_device(local)
dist = None
if world > 1:
    import torch.distributed as dist

    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
n = 3774873600  # one bench step of frames (4096 x 640 x 480 x 3)
h = torch.empty(n, dtype=torch.uint8).pin_memory()
d = torch.empty(n, dtype=torch.uint8, device=f"cuda:{local}")
for _ in range(2):
    d.copy_(h, non_blocking=True)
torch.cuda.synchronize()
if dist is not None:
    dist.barrier()
    torch.cuda.synchronize()
reps = 5
t0 = time.perf_counter()
for _ in range(reps):
    d.copy_(h, non_blocking=True)
torch.cuda.synchronize()
dt = time.perf_counter() - t0
if dist is not None:
    t = [torch.zeros(1, dtype=torch.float64, device="cuda") for _ in range(world)]
    dist.all_gather(t, torch.tensor([dt], dtype=torch.float64, device="cuda"))
    dts = [float(v.item()) for v in t]
else:
    dts = [dt]
if rank == 0:
    worst = max(dts)
    print(json.dumps({"n_gpus": world, "bytes_per_copy": n, "copies": reps,
                      "per_rank_gb_per_s": [n * reps / x / 1e9 for x in dts],
                      "aggregate_gb_per_s": world * n * reps / worst / 1e9,
                      "frames_per_s_ceiling_640x480": world * 4096 * reps / worst}))
if dist is not None:
    dist.barrier()
    dist.destroy_process_group()
