"""GPU parity of the densification kernel (fb_unproject_append) behind get_pointcloud / ResidentParams.append_from_frame
against (i) golden outputs of the reference's own get_pointcloud / initialize_new_params bodies (tests/golden/pcd_*.npz),
(ii) the numpy oracle, (iii) the reference's tensor expressions evaluated on the same GPU.  Colours, rotations, logits are
copies/constants: bit-exact.  Positions: rtol/atol 2e-6 (the 4-term dot product of c2w @ p, whose summation order inside
cuBLAS / numpy is not the kernel's FMA chain).  mean_sq_dist / log_scales: rtol 1e-6.
"""
import os

import numpy as np
import pytest

from tests.golden.make_golden import frame_inputs

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _t(torch, *xs):
    return [torch.from_numpy(np.ascontiguousarray(x)).cuda() for x in xs]


@pytest.mark.parametrize("name", ["pcd_a", "pcd_b", "pcd_c"])
def test_get_pointcloud_matches_reference_golden(cuda, oracle, name):
    import torch
    g = np.load(os.path.join(GOLD, name + ".npz"))
    H, W, ds = int(g["H"]), int(g["W"]), int(g["ds"])
    color, depth, K, w2c, mask = frame_inputs(H, W, int(g["seed"]))
    tc, td, tK, tw, tm = _t(torch, color, depth, K, w2c, mask)
    cld, msd = cuda.get_pointcloud(tc, td, tK, tw, mask=tm, compute_mean_sq_dist=True, downsample=ds)
    cld, msd = cld.cpu().numpy(), msd.cpu().numpy()
    assert cld.shape == g["cld"].shape and msd.shape == g["msd"].shape
    np.testing.assert_allclose(cld[:, :3], g["cld"][:, :3], rtol=2e-6, atol=2e-6)
    np.testing.assert_array_equal(cld[:, 3:], g["cld"][:, 3:])
    np.testing.assert_allclose(msd, g["msd"], rtol=1e-6)
    want, want_msd = oracle.get_pointcloud(color, depth, K, w2c, mask=mask, compute_mean_sq_dist=True, downsample=ds)
    np.testing.assert_allclose(cld, want, rtol=2e-6, atol=2e-6)
    np.testing.assert_array_equal(msd, want_msd)                          # same FP32 operations, same order
    full = cuda.get_pointcloud(tc, td, tK, tw, downsample=ds).cpu().numpy()
    np.testing.assert_allclose(full, g["cld_all"], rtol=2e-6, atol=2e-6)
    cam = cuda.get_pointcloud(tc, td, tK, tw, transform_pts=False, downsample=ds).cpu().numpy()
    np.testing.assert_allclose(cam, g["cld_cam"], rtol=1e-6, atol=1e-7)
    # a mask with nothing left in it: the reference warns and returns every pixel (:131-138)
    none = cuda.get_pointcloud(tc, td, tK, tw, mask=torch.zeros_like(tm), downsample=ds).cpu().numpy()
    np.testing.assert_array_equal(none, full)


def test_against_the_reference_expressions_on_the_same_gpu(cuda):
    """get_pointcloud's tensor program (gaussian.py:83-137) written out with torch on the GPU, full 256x256 frame."""
    import torch
    H = W = 256
    color, depth, K, w2c, mask = frame_inputs(H, W, 7)
    tc, td, tK, tw, tm = _t(torch, color, depth, K, w2c, mask)
    CX, CY, FX, FY = tK[0][2], tK[1][2], tK[0][0], tK[1][1]
    xg, yg = torch.meshgrid(torch.arange(0, W).cuda().float(), torch.arange(0, H).cuda().float(), indexing="xy")
    xx, yy = ((xg - CX) / FX).reshape(-1), ((yg - CY) / FY).reshape(-1)
    z = td[0].reshape(-1)
    pts4 = torch.cat((torch.stack((xx * z, yy * z, z), dim=-1), torch.ones(H * W, 1).cuda()), dim=1)
    pts = (torch.inverse(tw) @ pts4.T).T[:, :3]
    sg = 1 * z / ((FX + FY) / 2)
    want = torch.cat((pts, torch.permute(tc, (1, 2, 0)).reshape(-1, 3)), -1)[tm]
    want_msd = (sg ** 2)[tm]
    got, got_msd = cuda.get_pointcloud(tc, td, tK, tw, mask=tm, compute_mean_sq_dist=True)
    np.testing.assert_allclose(got.cpu().numpy(), want.cpu().numpy(), rtol=2e-6, atol=2e-6)
    assert torch.equal(got_msd, want_msd)
    assert torch.equal(got[:, 3:], want[:, 3:])


def test_resident_params_append_in_place(cuda, oracle):
    """append_from_frame == get_pointcloud + initialize_new_params + torch.cat (gaussian.py:353-355, 397-401), without
    re-allocating: the live rows keep their storage until capacity runs out, then the store grows once."""
    import torch
    H, W = 48, 64
    store = cuda.ResidentParams(capacity=1500, isotropic=False)
    want = {k: [] for k, _ in store.KEYS}
    ptr0 = store.buf["means3D"].data_ptr()
    for step, seed in enumerate((11, 12, 13)):
        color, depth, K, w2c, mask = frame_inputs(H, W, seed)
        mask &= depth.reshape(-1) > 0.01                                   # valid_depth_mask, :343-344
        r0, n = store.append_from_frame(*_t(torch, color, depth, K, w2c, mask), time_idx=step)
        cld, msd = oracle.get_pointcloud(color, depth, K, w2c, mask=mask, compute_mean_sq_dist=True)
        new = oracle.initialize_new_params(cld, msd)
        assert n == cld.shape[0] and r0 == sum(len(x) for x in want["means3D"])
        for k in want:
            want[k].append(new[k])
        if step == 0:
            assert store.buf["means3D"].data_ptr() == ptr0                 # written in place
    assert store.capacity > 1500 and store.buf["means3D"].data_ptr() != ptr0  # grew once the third frame did not fit
    p = store.params()
    for k in want:
        w = np.concatenate(want[k], axis=0)
        assert tuple(p[k].shape) == w.shape, k
        tol = dict(rtol=2e-6, atol=2e-6) if k in ("means3D", "log_scales") else dict(rtol=0, atol=0)
        np.testing.assert_allclose(p[k].cpu().numpy(), w, err_msg=k, **tol)
    ts = store.timestep[:store.n].cpu().numpy()
    assert ts[0] == 0 and ts[-1] == 2 and np.all(np.diff(ts) >= 0)
    # isotropic: one log-scale column (:309)
    iso = cuda.ResidentParams(capacity=10, isotropic=True)
    color, depth, K, w2c, mask = frame_inputs(H, W, 11)
    iso.append_from_frame(*_t(torch, color, depth, K, w2c, mask & (depth.reshape(-1) > 0.01)))
    assert iso.params()["log_scales"].shape[1] == 1
    np.testing.assert_array_equal(iso.params()["log_scales"][:, 0].cpu().numpy(), p["log_scales"][:iso.n, 0].cpu().numpy())
    # the appended rows feed the map the scorer consumes
    gmap = cuda.GaussianMap.from_params(store.params())
    assert gmap.P == store.n


def test_argument_checks(cuda):
    import torch
    color, depth, K, w2c, mask = frame_inputs(48, 64, 3)
    tc, td, tK, tw, tm = _t(torch, color, depth, K, w2c, mask)
    with pytest.raises(ValueError):
        cuda.get_pointcloud(tc, td, tK, tw, compute_mean_sq_dist=True, mean_sq_dist_method="knn")   # :117
    with pytest.raises(RuntimeError):
        cuda.get_pointcloud(tc.cpu(), td.cpu(), tK, tw)
    with pytest.raises(RuntimeError):
        cuda.get_pointcloud(tc, td, tK, tw, downsample=5)


def test_fused_rgb_depth_silhouette_render_matches_two_reference_calls(cuda):
    """SURVEY 8(f) N1, optimisation half: the two renders of get_loss (models/SLAM/gaussian.py:207,211) in ONE pass --
    image, depth + silhouette image, radii and every parameter gradient (through params AND transformed_pts, including
    the depth colours' dependence on the points) against two forward + two backward calls of the UNMODIFIED reference
    extension; means2D.grad is the colour render's alone."""
    import torch
    import torch.nn.functional as F
    from oracle import build_ref
    from fisher_nerf_customized_b200 import synth
    ref_C = build_ref.import_ref("ref_dgr")
    if ref_C is None:
        pytest.skip("oracle/_ref/ref_dgr not built")
    W = H = 256
    P = 200_000
    raw = synth.make_scene(P, seed=21)
    rs = cuda.setup_camera(W, H, synth.intrinsics(W, H), np.eye(4, dtype=np.float32))
    w2c = torch.from_numpy(synth.w2c_from_c2w(synth.make_poses(1, seed=4))[0]).cuda()
    g = torch.Generator(device="cuda").manual_seed(3)
    w_im = torch.randn((3, H, W), generator=g, device="cuda") * 1e-2
    w_ds = torch.randn((3, H, W), generator=g, device="cuda") * 1e-2
    e = torch.Tensor([])

    class RefRaster(torch.autograd.Function):        # the reference's autograd bridge (__init__.py:42-138), power 1
        @staticmethod
        def forward(ctx, tag, means3D, colors, opacities, scales, rotations):
            R, color, radii, geom, binning, img, depth = ref_C.rasterize_gaussians(
                rs.bg, means3D, colors, opacities, scales, rotations, 1.0, e, rs.viewmatrix, rs.projmatrix, rs.tanfovx,
                rs.tanfovy, H, W, e, 0, rs.campos, False)
            ctx.R, ctx.tag = R, tag
            ctx.save_for_backward(means3D, colors, scales, rotations, radii, geom, binning, img)
            return color, radii

        @staticmethod
        def backward(ctx, gc, _gr):
            means3D, colors, scales, rotations, radii, geom, binning, img = ctx.saved_tensors
            g2d, gcol, gop, g3d, gcov, gsh, gsc, grot = ref_C.rasterize_gaussians_backward(
                rs.bg, means3D, radii, colors, scales, rotations, 1.0, e, rs.viewmatrix, rs.projmatrix, rs.tanfovx,
                rs.tanfovy, gc.contiguous(), e, 0, rs.campos, geom, ctx.R, binning, img, 1)
            RefRaster.means2D[ctx.tag] = g2d
            return None, g3d, gcol, gop, gsc, grot

    def leaves():
        p = {k: torch.from_numpy(np.asarray(v, np.float32)).cuda().requires_grad_(True) for k, v in raw.items()}
        pts4 = torch.cat((p["means3D"], torch.ones(P, 1, device="cuda")), dim=1)
        return p, (w2c @ pts4.T).T[:, :3].contiguous()

    # reference: two renders
    p, tp = leaves()
    rot, op, sc = F.normalize(p["unnorm_rotations"]), torch.sigmoid(p["logit_opacities"]), torch.exp(p["log_scales"])
    RefRaster.means2D = {}
    im_r, radii_r = RefRaster.apply("rgb", tp, p["rgb_colors"], op, sc, rot)
    ds_r, _ = RefRaster.apply("ds", tp, cuda.get_depth_and_silhouette(tp, torch.eye(4, device="cuda")), op, sc, rot)
    ((im_r * w_im).sum() + (ds_r * w_ds).sum()).backward()
    m2d_r = RefRaster.means2D["rgb"]
    want = {k: v.grad.clone() for k, v in p.items()}
    # fused
    p2, tp2 = leaves()
    im, radius, depth_sil, means2D = cuda.render_rgb_depth_silhouette(p2, rs, torch.eye(4, device="cuda"), tp2)
    ((im * w_im).sum() + (depth_sil * w_ds).sum()).backward()
    assert torch.equal(radius, radii_r)
    assert torch.equal(im, im_r)                                  # same sweep arithmetic as the drop-in: bit-identical
    torch.testing.assert_close(depth_sil, ds_r, rtol=1e-5, atol=1e-6)
    vis = radius > 0
    for k in want:
        got, ref = p2[k].grad, want[k]
        scale = ref.abs().amax()
        bad = (got - ref).abs() > 1e-3 * ref.abs() + 1e-5 * scale
        assert bad[vis].float().mean().item() < 1e-3, (k, int(bad.sum()))
        assert not got[~vis].any()
    scale = m2d_r.abs().amax()
    bad = (means2D.grad - m2d_r).abs() > 1e-3 * m2d_r.abs() + 1e-5 * scale
    assert bad.float().mean().item() < 1e-3
