"""fb_occ_update_static -- the occupancy update with its scratch in a file-scope device array (no cudaMalloc, no caller
workspace; include/fisher_b200.h) -- against fb_occ_update, called through the C ABI on the same inputs, two calls in a
row.  The per-cell counts are integers, so the two maps must be bit-identical (the tolerance of
tests/test_gpu_mapnbr.py).  This file is named to be collected last."""
import ctypes

import numpy as np
import pytest

from tests import common

pytestmark = pytest.mark.gpu

CELL, H_LO, H_HI, FAR = 0.05, -0.6, 0.8, 5.0
FB_E_STATIC_LIMIT = -6
MAX_CELLS = 1024 * 1024


def _depth(H, W, K, c2w, rng):
    """z-depth of a 6 x 3 x 6 m box room seen from c2w, with holes (0) and far pixels."""
    ys, xs = np.mgrid[0:H, 0:W].astype(np.float32)
    d = np.stack([(xs - K[0, 2]) / K[0, 0], (ys - K[1, 2]) / K[1, 1], np.ones_like(xs)], -1) @ c2w[:3, :3].T
    o = c2w[:3, 3]
    half = np.array([3.0, 1.5, 3.0])
    with np.errstate(divide="ignore", invalid="ignore"):
        t = np.where(d > 0, (half - o) / d, (-half - o) / d)
    depth = np.nanmin(t, axis=-1).astype(np.float32)
    depth[rng.random((H, W)) < 0.05] = 0.0
    depth[rng.random((H, W)) < 0.02] = 50.0
    return depth


def test_static_scratch_entry_point_matches_workspace_entry_point(cuda):
    import torch
    L = cuda._lib.lib()
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    rng = np.random.default_rng(11)
    H, W = 60, 80
    K = np.array([[60.0, 0, 40.0], [0, 60.0, 30.0], [0, 0, 1]], np.float32)
    intr = torch.tensor([K[0, 0], K[1, 1], K[0, 2], K[1, 2]], device=dev)
    planes = torch.linspace(1e-3, 0.95, 11, device=dev)
    planes[-1] = 1.0
    centre = torch.zeros(2, device=dev)
    poses = [np.linalg.inv(common.look_at_w2c((0.2, 0.0, -0.4), (2.0, 0.1, 2.0))).astype(np.float32),
             np.linalg.inv(common.look_at_w2c((0.9, 0.05, 0.3), (-2.0, 0.0, 1.0))).astype(np.float32)]
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)

    def run(static, occ, gx, gz, depth, c2w, cam, ds):
        if static:
            return L.fb_occ_update_static(occ.data_ptr(), gx, gz, depth.data_ptr(), W, H, ds, c2w.data_ptr(), intr.data_ptr(),
                                          planes.data_ptr(), centre.data_ptr(), CELL, H_LO, H_HI, FAR, cam.data_ptr(), stream)
        ws = torch.empty((L.fb_occ_workspace_bytes(gx, gz),), dtype=torch.uint8, device=dev)
        return L.fb_occ_update(occ.data_ptr(), gx, gz, depth.data_ptr(), W, H, ds, c2w.data_ptr(), intr.data_ptr(),
                               planes.data_ptr(), centre.data_ptr(), CELL, H_LO, H_HI, FAR, cam.data_ptr(), ws.data_ptr(),
                               ws.numel(), stream)

    # a large grid first (the reference's 768 x 768), then a small one: what the first call left in the shared scratch
    # must not leak into the second; then the large one again, two calls in a row on the same map
    for gx, gz, ds in ((768, 768, 1), (141, 141, 2), (768, 768, 1)):
        occ_a = torch.zeros((3, gz, gx), device=dev)
        occ_a[0] = 1.0
        occ_b = occ_a.clone()
        cam_a = torch.zeros(2, dtype=torch.int32, device=dev)
        cam_b = torch.zeros(2, dtype=torch.int32, device=dev)
        for call, c2w_np in enumerate(poses):
            depth = torch.from_numpy(_depth(H, W, K, c2w_np, rng)).to(dev).contiguous()
            c2w = torch.from_numpy(c2w_np).to(dev).contiguous()
            assert run(False, occ_a, gx, gz, depth, c2w, cam_a, ds) == 0
            torch.cuda.synchronize()
            held, free0 = torch.cuda.memory_allocated(), torch.cuda.mem_get_info()[0]
            rc = run(True, occ_b, gx, gz, depth, c2w, cam_b, ds)
            torch.cuda.synchronize()
            assert rc == 0, cuda._lib.lib().fb_error_string(rc)
            assert torch.cuda.memory_allocated() == held                     # nothing from torch's allocator
            if call == 1:                                                    # (the first call may load the module lazily)
                assert torch.cuda.mem_get_info()[0] >= free0 - (2 << 20)      # nothing from the driver's either
            assert torch.equal(cam_a, cam_b)
            assert torch.equal(occ_a, occ_b), f"grid {gx}x{gz}, call {call}: {int((occ_a != occ_b).sum())} cells differ"
        got = occ_b.cpu().numpy()
        assert (got[1] > 0).sum() > 50 and (got[2] > 0).sum() > 500          # the update did something

    # beyond the static maximum: a clear error, nothing touched
    gx, gz = 2048, 1024
    assert gx * gz > MAX_CELLS
    occ = torch.zeros((3, 4, 4), device=dev)            # never dereferenced: the size check comes first
    before = occ.clone()
    rc = run(True, occ, gx, gz, depth, c2w, cam_b, 1)
    assert rc == FB_E_STATIC_LIMIT
    assert b"static scratch" in L.fb_error_string(rc)
    torch.cuda.synchronize()
    assert torch.equal(occ, before)
    # exactly at the maximum is accepted
    occ = torch.zeros((3, 1024, 1024), device=dev)
    assert run(True, occ, 1024, 1024, depth, c2w, cam_b, 1) == 0
    torch.cuda.synchronize()
