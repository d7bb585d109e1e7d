"""GPU parity of the drop-in rasteriser (stages 1-4, single view) through the C ABI:
CUDA path vs the CPU oracle, and vs the reference CUDA extension rebuilt for sm_100a (oracle/_ref)
when it travelled with the snapshot.

Tolerances (BASELINE.json north_star): bit-exact radii / tile keys / sorted list / ranges / n_contrib;
rtol 1e-4 images; rtol 1e-3 gradients and Hessian diagonals.
"""
import numpy as np
import pytest

from tests import common

pytestmark = pytest.mark.gpu

IMG_RTOL, IMG_ATOL = 1e-4, 1e-6
H_RTOL = 1e-3


def _settings(fb, torch, W, H, w2c=None):
    return fb.setup_camera(W, H, common.K_for(W, H), np.eye(4, dtype=np.float32) if w2c is None else w2c)


def _to_t(torch, sc):
    return {k: torch.from_numpy(v).cuda() for k, v in sc.items()}


def _run_cuda(fb, torch, rs, t, power, dpix, want_debug=True):
    P = t["means3D"].shape[0]
    args = (rs.bg, t["means3D"], t["rgb_colors"], t["opacities"], t["scales"], t["rotations"], rs.scale_modifier,
            torch.Tensor([]), rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, rs.image_height, rs.image_width,
            torch.Tensor([]), rs.sh_degree, rs.campos, rs.prefiltered)
    R, color, radii, geom, binning, img, depth = fb._C.rasterize_gaussians(*args)
    out = dict(R=R, color=color.cpu().numpy(), radii=radii.cpu().numpy(), depth=depth.cpu().numpy())
    if want_debug:
        out.update(fb.rasterizer.debug_export(rs, P, R, geom, binning, img))
    if power:
        g = fb._C.rasterize_gaussians_backward(rs.bg, t["means3D"], radii, t["rgb_colors"], t["scales"], t["rotations"],
                                                rs.scale_modifier, torch.Tensor([]), rs.viewmatrix, rs.projmatrix,
                                                rs.tanfovx, rs.tanfovy, dpix, torch.Tensor([]), rs.sh_degree, rs.campos,
                                                geom, R, binning, img, power)
        names = ("dL_dmeans2D", "dL_dcolors", "dL_dopacity", "dL_dmeans3D", "dL_dcov3D", "dL_dsh", "dL_dscales",
                 "dL_drotations")
        out.update({n: v.cpu().numpy() for n, v in zip(names, g)})
    return out


def _run_oracle(O, rs, sc):
    cam = O.make_camera(rs.image_width, rs.image_height, rs.tanfovx, rs.tanfovy, rs.viewmatrix.cpu().numpy(),
                        rs.projmatrix.cpu().numpy(), rs.bg.cpu().numpy(), rs.scale_modifier, rs.sh_degree, 0,
                        rs.campos.cpu().numpy(), rs.prefiltered)
    return O.Forward(cam, sc["means3D"], sc["opacities"], colors_precomp=sc["rgb_colors"], scales=sc["scales"],
                     rotations=sc["rotations"])


def _close(a, b, rtol, atol):
    return np.abs(a - b) <= atol + rtol * np.abs(b)


def _assert_grads(got, want, scale_ref, rtol=H_RTOL, frac=1e-4):
    """Per-tensor check: |got-want| <= rtol*|want| + rtol*1e-3*max|want| for all but a vanishing
    fraction of entries (pairs whose alpha sits on the 1/255 threshold flip with the exp() ulp)."""
    atol = rtol * 1e-3 * max(scale_ref, 1e-30)
    bad = ~_close(got.astype(np.float64), want.astype(np.float64), rtol, atol)
    assert bad.mean() <= frac, f"{bad.sum()} of {bad.size} entries off (max rel " \
                               f"{np.max(np.abs(got - want) / (np.abs(want) + atol)):.3g})"


@pytest.mark.parametrize("P,W,H,seed", [(3000, 64, 64, 1), (20000, 256, 256, 2), (5000, 100, 70, 3)])
def test_forward_matches_oracle(cuda, oracle, P, W, H, seed):
    import torch
    sc = common.random_view_scene(P, seed, W, H)
    rs = _settings(cuda, torch, W, H)
    got = _run_cuda(cuda, torch, rs, _to_t(torch, sc), 0, None)
    ref = _run_oracle(oracle, rs, sc)
    # integer / index outputs: bit-exact
    assert got["R"] == ref.num_rendered
    np.testing.assert_array_equal(got["radii"], ref.radii)
    np.testing.assert_array_equal(got["tiles_touched"], ref.tiles_touched)
    np.testing.assert_array_equal(got["ranges"], ref.ranges)
    np.testing.assert_array_equal(got["point_keys"], ref.keys)
    np.testing.assert_array_equal(got["point_list"], ref.point_list)
    vis = ref.radii > 0
    np.testing.assert_array_equal(got["depths"][vis], ref.depths[vis])
    np.testing.assert_array_equal(got["means2D"][vis], ref.means2D[vis])
    np.testing.assert_array_equal(got["conic_opacity"][vis], ref.conic_opacity[vis])
    # n_contrib: exact up to the ulp of exp() (CUDA expf vs glibc): allow a vanishing fraction
    nc_bad = (got["n_contrib"] != ref.n_contrib).mean()
    assert nc_bad <= 2e-3, nc_bad
    ok = _close(got["color"], ref.color, IMG_RTOL, IMG_ATOL)
    assert ok.mean() >= 1 - 2e-3, f"colour mismatch fraction {1 - ok.mean():.3g}"
    assert np.abs(got["color"] - ref.color).max() < 2e-2
    ok = _close(got["final_T"], ref.final_T, 1e-4, 1e-7)
    assert ok.mean() >= 1 - 2e-3
    assert (got["depth"] == ref.depth).mean() >= 1 - 2e-3   # median depth is a selection


@pytest.mark.parametrize("power", [1, 2])
@pytest.mark.parametrize("P,W,H,seed", [(3000, 64, 64, 11), (12000, 128, 128, 12)])
def test_backward_matches_oracle(cuda, oracle, power, P, W, H, seed):
    import torch
    sc = common.random_view_scene(P, seed, W, H)
    rs = _settings(cuda, torch, W, H)
    rng = np.random.default_rng(seed)
    dpix = (rng.normal(size=(3, H, W)) * 1e-3).astype(np.float32) if power == 1 else np.full((3, H, W), 1e-3, np.float32)
    got = _run_cuda(cuda, torch, rs, _to_t(torch, sc), power, torch.from_numpy(dpix).cuda(), want_debug=False)
    ref = _run_oracle(oracle, rs, sc)
    want = ref.backward(dpix, power)
    for name in ("dL_dmeans2D", "dL_dcolors", "dL_dopacity", "dL_dmeans3D", "dL_dcov3D", "dL_dscales", "dL_drotations"):
        g, w = got[name].reshape(P, -1), want[name].reshape(P, -1)
        if power == 1:
            # signed sums cancel: compare against the magnitude scale of the tensor
            scale = np.abs(w).max()
            err = np.abs(g - w).max()
            assert err <= 2e-3 * scale + 1e-12, (name, err, scale)
        else:
            _assert_grads(g, w, np.abs(w).max(), frac=2e-3)
    # invisible Gaussians have exactly zero in every column (rasterize_points.cu:151-159)
    inv = ref.radii == 0
    for name in ("dL_dmeans3D", "dL_dopacity", "dL_dscales", "dL_drotations"):
        assert not got[name].reshape(P, -1)[inv].any()


def _ref_ext():
    from oracle import build_ref
    return build_ref.import_ref("ref_dgr")


@pytest.mark.parametrize("P,W,H,seed,ident", [(20000, 256, 256, 21, True), (8000, 128, 96, 22, False)])
def test_forward_backward_match_reference_extension(cuda, P, W, H, seed, ident):
    """Same inputs through the UNMODIFIED reference CUDA extension (oracle/_ref) on this GPU."""
    import torch
    ref_C = _ref_ext()
    if ref_C is None:
        pytest.skip("oracle/_ref/ref_dgr not built (built only where /root/reference exists)")
    sc = common.random_view_scene(P, seed, W, H)
    w2c = None if ident else common.look_at_w2c((0.3, -0.2, -0.5), (0.1, 0.0, 3.0))
    rs = _settings(cuda, torch, W, H, w2c)
    t = _to_t(torch, sc)
    dpix = torch.full((3, H, W), 1e-3, device="cuda")
    for power in (1, 2):
        got = _run_cuda(cuda, torch, rs, t, power, dpix)
        e = torch.Tensor([])
        R, color, radii, geom, binning, img, depth = ref_C.rasterize_gaussians(
            rs.bg, t["means3D"], t["rgb_colors"], t["opacities"], t["scales"], t["rotations"], rs.scale_modifier, e,
            rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, rs.image_height, rs.image_width, e, rs.sh_degree,
            rs.campos, rs.prefiltered)
        assert got["R"] == R
        np.testing.assert_array_equal(got["radii"], radii.cpu().numpy())
        from tests.golden.make_golden import _img_state
        ref_T, ref_nc, ref_ranges = _img_state(img, H * W)
        tiles = got["ranges"].shape[0]
        np.testing.assert_array_equal(got["ranges"], ref_ranges[:2 * tiles].reshape(tiles, 2))
        np.testing.assert_array_equal(got["n_contrib"], ref_nc)
        # the reference's own GeometryState / BinningState (rasterizer_impl.cu:155-194): the sorted instance list and
        # its (tile | depth) keys, tiles_touched and the per-Gaussian projection, bit for bit
        from tests.golden.make_golden import _geom_state, _binning_state
        r_depths, r_means2D, _, r_conic_op, r_tiles = _geom_state(geom, P)
        r_list, r_keys = _binning_state(binning, R)
        vis = radii.cpu().numpy() > 0
        np.testing.assert_array_equal(got["point_list"], r_list)
        np.testing.assert_array_equal(got["point_keys"].view(np.uint64), r_keys)
        np.testing.assert_array_equal(got["tiles_touched"][vis], r_tiles[vis])
        np.testing.assert_array_equal(got["depths"][vis], r_depths[vis])
        np.testing.assert_array_equal(got["means2D"][vis], r_means2D[vis])
        np.testing.assert_array_equal(got["conic_opacity"][vis], r_conic_op[vis])
        np.testing.assert_allclose(got["final_T"], ref_T, rtol=1e-5, atol=1e-8)
        np.testing.assert_allclose(got["color"], color.cpu().numpy(), rtol=IMG_RTOL, atol=IMG_ATOL)
        np.testing.assert_array_equal(got["depth"], depth.cpu().numpy())
        g = ref_C.rasterize_gaussians_backward(rs.bg, t["means3D"], radii, t["rgb_colors"], t["scales"], t["rotations"],
                                               rs.scale_modifier, e, rs.viewmatrix, rs.projmatrix, rs.tanfovx,
                                               rs.tanfovy, dpix, e, rs.sh_degree, rs.campos, geom, R, binning, img, power)
        names = ("dL_dmeans2D", "dL_dcolors", "dL_dopacity", "dL_dmeans3D", "dL_dcov3D", "dL_dsh", "dL_dscales",
                 "dL_drotations")
        for n, w in zip(names, g):
            if n == "dL_dsh":
                continue
            w = w.cpu().numpy().reshape(P, -1)
            gg = got[n].reshape(P, -1)
            if power == 1:
                scale = np.abs(w).max()
                assert np.abs(gg - w).max() <= 2e-3 * scale + 1e-12, n
            else:
                # ~10 of 8000 Gaussians own a pair whose alpha sits on the 1/255 threshold or whose atomics land in a
                # different order run to run (observed 0.6e-3 .. 1.04e-3 of the entries): bound at 2e-3
                _assert_grads(gg, w, np.abs(w).max(), frac=2e-3)


def test_operator_api_autograd(cuda):
    """GaussianRasterizer drop-in: shapes, dtypes, autograd wiring, argument validation."""
    import torch
    W = H = 64
    sc = common.random_view_scene(2000, 5, W, H)
    t = _to_t(torch, sc)
    rs = _settings(cuda, torch, W, H)
    for k in t:
        t[k].requires_grad_(True)
    means2D = torch.zeros_like(t["means3D"], requires_grad=True)
    rast = cuda.GaussianRasterizer(raster_settings=rs, backward_power=2)
    im, radii, depth = rast(means3D=t["means3D"], means2D=means2D, opacities=t["opacities"],
                            colors_precomp=t["rgb_colors"], scales=t["scales"], rotations=t["rotations"])
    assert im.shape == (3, H, W) and radii.shape == (2000,) and radii.dtype == torch.int32 and depth.shape == (1, H, W)
    im.backward(gradient=torch.ones_like(im) * 1e-3)
    assert t["means3D"].grad.shape == (2000, 3) and t["opacities"].grad.shape == (2000, 1)
    assert (t["means3D"].grad >= 0).all() and t["means3D"].grad.sum() > 0    # power 2: sums of squares
    with pytest.raises(Exception):
        rast(means3D=t["means3D"], means2D=means2D, opacities=t["opacities"], scales=t["scales"], rotations=t["rotations"])
    with pytest.raises(Exception):
        rast(means3D=t["means3D"], means2D=means2D, opacities=t["opacities"], colors_precomp=t["rgb_colors"])
    vis = rast.markVisible(t["means3D"].detach())
    np.testing.assert_array_equal(vis.cpu().numpy(), sc["means3D"][:, 2] > 0.001)


def test_empty_and_culled_inputs(cuda):
    import torch
    W = H = 32
    rs = _settings(cuda, torch, W, H)
    e = torch.Tensor([])
    z = torch.zeros((0, 3), device="cuda")
    R, color, radii, *_ = cuda._C.rasterize_gaussians(rs.bg, z, z, torch.zeros((0, 1), device="cuda"), z,
                                                      torch.zeros((0, 4), device="cuda"), 1.0, e, rs.viewmatrix,
                                                      rs.projmatrix, rs.tanfovx, rs.tanfovy, H, W, e, 0, rs.campos, False)
    assert R == 0 and color.shape == (3, H, W) and not color.any()
    # everything behind the camera: nothing rendered, background only
    sc = common.random_view_scene(500, 9, W, H)
    sc["means3D"][:, 2] = -np.abs(sc["means3D"][:, 2]) - 0.1
    t = _to_t(torch, sc)
    got = _run_cuda(cuda, torch, rs, t, 2, torch.full((3, H, W), 1e-3, device="cuda"), want_debug=False)
    assert got["R"] == 0 and not got["radii"].any() and not got["color"].any()
    assert not got["dL_dmeans3D"].any()
    with pytest.raises(RuntimeError):
        cuda._C.rasterize_gaussians(rs.bg, torch.zeros((4, 2), device="cuda"), z, z, z, z, 1.0, e, rs.viewmatrix,
                                    rs.projmatrix, rs.tanfovx, rs.tanfovy, H, W, e, 0, rs.campos, False)


def test_long_tile_lists_and_depth_ties(cuda, oracle):
    """Tile lists longer than the shared-memory sort capacity (4096 keys) and thousands of bit-identical
    depths: the sorted (key, id) list must still equal the reference order (stable radix sort of keys
    emitted in ascending Gaussian index)."""
    import torch
    W = H = 32
    P = 9000
    sc = common.random_view_scene(P, 31, W, H)
    rng = np.random.default_rng(31)
    sc["means3D"][:, :2] = rng.uniform(-0.3, 0.3, (P, 2)).astype(np.float32)
    sc["means3D"][:, 2] = np.round(rng.uniform(1.0, 3.0, P), 1).astype(np.float32)   # 21 distinct depths
    sc["scales"][:] = 0.02
    rs = _settings(cuda, torch, W, H)
    got = _run_cuda(cuda, torch, rs, _to_t(torch, sc), 0, None)
    ref = _run_oracle(oracle, rs, sc)
    assert (ref.ranges[:, 1] - ref.ranges[:, 0]).max() > 4096
    np.testing.assert_array_equal(got["ranges"], ref.ranges)
    np.testing.assert_array_equal(got["point_keys"], ref.keys)
    np.testing.assert_array_equal(got["point_list"], ref.point_list)
    assert (got["n_contrib"] != ref.n_contrib).mean() <= 2e-3
