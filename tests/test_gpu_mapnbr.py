"""GPU parity of the occupancy update and the masked visibility counter against the numpy oracle
(oracle.update_occ_map / oracle.count_gaussians_vs_mask) and the fixtures made from the reference's own function body
(tests/golden/mask_*.npz).  Integer results (per-cell counts through the fused map, index lists) are exact."""
import os

import numpy as np
import pytest

from tests import common

pytestmark = pytest.mark.gpu


def test_count_gaussians_vs_mask_matches_reference_body_and_oracle(cuda, oracle):
    import torch
    from tests.golden import make_golden as G
    for name, N, H, W, seed in G.MASK_CASES:
        means, logit, w2c, K, mask = G.mask_inputs(N, H, W, seed)
        z = np.load(os.path.join(os.path.dirname(G.__file__), name + ".npz"))
        params = {"means3D": torch.from_numpy(means).cuda(), "logit_opacities": torch.from_numpy(logit).cuda()}
        r = cuda.count_gaussians_vs_mask(params, torch.from_numpy(w2c).cuda(), torch.from_numpy(K), torch.from_numpy(mask), 0.01)
        assert (r["in"], r["out"], r["active"]) == (int(z["n_in"]), int(z["n_out"]), int(z["active"]))
        np.testing.assert_array_equal(r["idx_in"].cpu().numpy(), z["idx_in"])
        np.testing.assert_array_equal(r["idx_out"].cpu().numpy(), z["idx_out"])
        o = oracle.count_gaussians_vs_mask(means, logit, w2c, (K[0, 0], K[1, 1], K[0, 2], K[1, 2]), mask, 0.01)
        np.testing.assert_array_equal(r["idx_in"].cpu().numpy(), o["idx_in"])
    # edge cases of the reference: no mask, nothing active
    params = {"means3D": torch.zeros((5, 3), device="cuda"), "logit_opacities": torch.full((5, 1), -20.0, device="cuda")}
    r = cuda.count_gaussians_vs_mask(params, torch.eye(4), np.eye(3), None)
    assert (r["in"], r["out"], r["active"]) == (0, 5, 0)
    r = cuda.count_gaussians_vs_mask(params, torch.eye(4), np.eye(3), torch.ones((8, 8)))
    assert (r["in"], r["out"], r["active"]) == (0, 0, 0) and r["idx_in"].numel() == 0


def _room_depth(H, W, K, c2w, rng):
    """Depth of a 6 x 3 x 6 m box room seen from c2w (ray / box intersection), with holes and far pixels."""
    ys, xs = np.mgrid[0:H, 0:W].astype(np.float32)
    d = np.stack([(xs - K[0, 2]) / K[0, 0], (ys - K[1, 2]) / K[1, 1], np.ones_like(xs)], -1) @ c2w[:3, :3].T
    o = c2w[:3, 3]
    half = np.array([3.0, 1.5, 3.0])
    with np.errstate(divide="ignore", invalid="ignore"):
        t = np.where(d > 0, (half - o) / d, (-half - o) / d)
    depth = np.nanmin(t, axis=-1).astype(np.float32)       # z-depth: the ray direction has unit camera z
    depth[rng.random((H, W)) < 0.05] = 0.0
    depth[rng.random((H, W)) < 0.02] = 50.0
    return depth[None]


def test_update_occ_map_matches_oracle(cuda, oracle):
    import torch
    rng = np.random.default_rng(7)
    H, W = 60, 80
    K = np.array([[60.0, 0, 40.0], [0, 60.0, 30.0], [0, 0, 1]], np.float32)
    poses = [np.linalg.inv(common.look_at_w2c((0.2, 0.0, -0.4), (2.0, 0.1, 2.0))).astype(np.float32),
             np.linalg.inv(common.look_at_w2c((0.9, 0.05, 0.3), (-2.0, 0.0, 1.0))).astype(np.float32)]
    for scene_bounds, ds in ((None, 1), ((np.array([-3.5, -1.5, -3.5]), np.array([3.5, 1.5, 3.5])), 2)):
        m = cuda.OccupancyMap(cell_size=0.05, height_lower=-0.6, height_upper=0.8, pcd_far_distance=5.0)
        m.init(poses[0], K, scene_bounds)
        want = m.occ_map.cpu().numpy().copy()
        planes = m._planes.cpu().numpy()
        for t, c2w in enumerate(poses):
            depth = _room_depth(H, W, K, c2w, rng)
            m.update_occ_map(torch.from_numpy(depth).cuda(), torch.from_numpy(c2w).cuda(), t, downsample=ds)
            cam = oracle.update_occ_map(want, depth, c2w, (K[0, 0], K[1, 1], K[0, 2], K[1, 2]), 0.05, m.map_center.cpu().numpy(),
                                        -0.6, 0.8, 5.0, planes, downsample=ds)
            got = m.occ_map.cpu().numpy()
            assert tuple(m.cam_pos) == (cam[1], cam[0])
            # which cells were touched, and how (occupied / free / ray), is integer work: exact
            np.testing.assert_array_equal(got != 0, want != 0)
            np.testing.assert_array_equal(got, want)
        assert (got[1] > 0).sum() > 50 and (got[2] > 0).sum() > 500
