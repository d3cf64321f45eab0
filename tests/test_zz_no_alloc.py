"""GPU: mantis_edges_noalloc (scratch in a fixed __device__ array, no cudaMalloc inside the call) against
mantis_edges_batch on the same device-resident frames, through the C ABI: bit-exact (the tolerance of every
image-stage test here), on calls in a row whose shapes differ, so each one runs on the previous one's
leftovers.  Collected last on purpose."""
import numpy as np
import pytest

from cases import blocky

pytestmark = pytest.mark.gpu


def _both(ctx, frames):
    import torch

    n, h, w, _ = frames.shape
    dev = torch.device("cuda:0")
    d_in = torch.from_numpy(frames).to(dev)
    ref_c = torch.empty((n, h, w), dtype=torch.uint8, device=dev)
    ref_m = torch.empty_like(ref_c)
    new_c = torch.full_like(ref_c, 7)
    new_m = torch.full_like(ref_c, 7)
    ctx.edges_device(d_in.data_ptr(), n, w, h, d_canny=ref_c.data_ptr(), d_morph=ref_m.data_ptr())
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    ctx.edges_noalloc(d_in.data_ptr(), n, w, h, d_canny=new_c.data_ptr(), d_morph=new_m.data_ptr())
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    return ref_c.cpu().numpy(), ref_m.cpu().numpy(), new_c.cpu().numpy(), new_m.cpu().numpy(), free0, free1


def test_noalloc_matches_edges_batch_on_two_calls_in_a_row(ctx):
    import mantis_b200
    from oracle import pyoracle as po

    rng = np.random.default_rng(0x4D414E54 + 20)
    # call 1: dense edges, the marching kernel's shape (W % 16 == 0), near the scratch limit (100 of 109 frames)
    a = np.stack([blocky(rng, 480, 640, 9 + 3 * (i % 7)) for i in range(100)])
    # call 2: another shape (the tiled kernel: W % 16 != 0, partial last word), sparse edges, fewer frames
    b = np.stack([blocky(rng, 123, 77, 40) for _ in range(5)])
    for k, frames in enumerate((a, b, a[:3], b[:1])):
        rc, rm, nc, nm, free0, free1 = _both(ctx, frames)
        assert np.array_equal(nc, rc), f"call {k}: canny differs at {np.argwhere(nc != rc)[:5]}"
        assert np.array_equal(nm, rm), f"call {k}: morph differs at {np.argwhere(nm != rm)[:5]}"
        if k > 0:  # (the first call may load the module that holds the array)
            assert free1 >= free0, f"call {k}: device memory shrank by {free0 - free1} bytes inside the call"
    oc, om = po.edge_stage(b[0], 50)  # and both equal the oracle
    assert np.array_equal(nc[0], oc) and np.array_equal(nm[0], om)
    assert mantis_b200.binding.NOALLOC_SCRATCH_BYTES == 8 << 20


def test_noalloc_rejects_what_does_not_fit(ctx):
    import torch

    import mantis_b200

    n, h, w = 110, 480, 640  # 110 x 38 400 B per plane > 4 MiB
    d_in = torch.zeros((n, h, w, 3), dtype=torch.uint8, device="cuda:0")
    out = torch.empty((n, h, w), dtype=torch.uint8, device="cuda:0")
    with pytest.raises(mantis_b200.MantisError, match="MANTIS_NOALLOC_SCRATCH_BYTES"):
        ctx.edges_noalloc(d_in.data_ptr(), n, w, h, d_canny=out.data_ptr())
    with pytest.raises(mantis_b200.MantisError):  # no output at all
        ctx.edges_noalloc(d_in.data_ptr(), 1, w, h)
    ctx.edges_noalloc(d_in.data_ptr(), 109, w, h, d_canny=out.data_ptr())  # the largest batch that fits
    torch.cuda.synchronize()
    assert int(out[:109].max()) == 0  # a blank frame has no edges
