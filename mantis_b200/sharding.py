"""Frame sharding across the GPUs of one box: frames are independent, so GPU g of G
gets the contiguous range [g*N/G, (g+1)*N/G) and no collective runs on the data path.  Per-frame
results (88-byte records) are gathered to rank 0 over the host-side process group afterwards."""
import ctypes

import numpy as np

from .binding import Result


def frame_shard(n_frames, world, rank):
    """Contiguous shard of `n_frames` for `rank` of `world`: (first, count).  Sizes differ by at most one."""
    if world <= 0 or not 0 <= rank < world:
        raise ValueError("bad world / rank")
    base, extra = divmod(n_frames, world)
    first = rank * base + min(rank, extra)
    return first, base + (1 if rank < extra else 0)


def results_to_array(results):
    """ctypes array of mantis_result -> uint8 numpy array [n, sizeof(mantis_result)] (a copy)."""
    n = len(results)
    raw = np.frombuffer(bytes(memoryview(results)), dtype=np.uint8) if n else np.zeros(0, np.uint8)
    return raw.reshape(n, ctypes.sizeof(Result)).copy()


def array_to_results(arr):
    arr = np.ascontiguousarray(arr, np.uint8)
    n = arr.shape[0]
    out = (Result * n)()
    if n:
        ctypes.memmove(ctypes.addressof(out), arr.ctypes.data, arr.nbytes)
    return out


def gather_results(results, dist=None, n_total=None):
    """Gathers every rank's shard of results on rank 0, in frame order (shards are contiguous and
    ordered by rank).  `dist` is an initialised torch.distributed module or None (single process).
    Returns the full ctypes array on rank 0 and None elsewhere."""
    local = results_to_array(results)
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return array_to_results(local)
    import torch

    world, rank = dist.get_world_size(), dist.get_rank()
    rec = ctypes.sizeof(Result)
    counts = [torch.zeros(1, dtype=torch.int64) for _ in range(world)]
    dist.all_gather(counts, torch.tensor([local.shape[0]], dtype=torch.int64))
    counts = [int(c.item()) for c in counts]
    cap = max(counts) if counts else 0
    padded = torch.zeros((cap, rec), dtype=torch.uint8)
    if local.shape[0]:
        padded[: local.shape[0]] = torch.from_numpy(local)
    bufs = [torch.zeros((cap, rec), dtype=torch.uint8) for _ in range(world)] if rank == 0 else None
    dist.gather(padded, bufs, dst=0)
    if rank != 0:
        return None
    full = np.concatenate([bufs[r][: counts[r]].numpy() for r in range(world)], axis=0)
    if n_total is not None and full.shape[0] != n_total:
        raise RuntimeError(f"gathered {full.shape[0]} results, expected {n_total}")
    return array_to_results(full)
