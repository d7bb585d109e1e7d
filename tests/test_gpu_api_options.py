"""GPU parity of the operator-API options no Fisher caller uses but the drop-in must honour (SURVEY 8(a) note):
SH colours of degree 0-3, cov3D_precomp, non-zero background, scale_modifier != 1, prefiltered, and the
reference's argument checks (DGR/diff_gaussian_rasterization/__init__.py:174-178).  Checked against the UNMODIFIED
reference extension on this GPU (oracle/_ref) where it travelled, else against the CPU oracle.
"""
import numpy as np
import pytest

from tests import common

pytestmark = pytest.mark.gpu


def _ref_ext():
    from oracle import build_ref
    return build_ref.import_ref("ref_dgr")


def _t(torch, sc):
    return {k: torch.from_numpy(v).cuda() for k, v in sc.items()}


def _fwd(C, torch, rs, t, shs=None, cov3D=None, degree=0):
    e = torch.Tensor([])
    return C.rasterize_gaussians(
        rs.bg, t["means3D"], e if shs is not None else t["rgb_colors"], t["opacities"],
        e if cov3D is not None else t["scales"], e if cov3D is not None else t["rotations"], rs.scale_modifier,
        cov3D if cov3D is not None else e, rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, rs.image_height,
        rs.image_width, shs if shs is not None else e, degree, rs.campos, rs.prefiltered)


def _bwd(C, torch, rs, t, fw, dpix, power, shs=None, cov3D=None, degree=0):
    e = torch.Tensor([])
    R, color, radii, geom, binning, img, depth = fw
    return C.rasterize_gaussians_backward(
        rs.bg, t["means3D"], radii, e if shs is not None else t["rgb_colors"],
        e if cov3D is not None else t["scales"], e if cov3D is not None else t["rotations"], rs.scale_modifier,
        cov3D if cov3D is not None else e, rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, dpix,
        shs if shs is not None else e, degree, rs.campos, geom, R, binning, img, power)


def _cov3d_np(scales, q, mod=1.0):
    """Sigma = (S R)^T (S R) packed as the reference packs it (forward.cu:118-152), float64 then rounded."""
    r, x, y, z = (q[:, i].astype(np.float64) for i in range(4))
    R = np.stack([1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y),
                  2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
                  2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], axis=1).reshape(-1, 3, 3)
    S = (scales.astype(np.float64) * mod)[:, :, None] * np.eye(3)[None]
    M = R @ S                                # glm: M = S * R with column-major R == (R S)^T in row-major terms
    Sig = M @ M.transpose(0, 2, 1)
    return np.stack([Sig[:, 0, 0], Sig[:, 0, 1], Sig[:, 0, 2], Sig[:, 1, 1], Sig[:, 1, 2], Sig[:, 2, 2]], 1).astype(np.float32)


def _settings(fb, torch, W, H, bg=(0, 0, 0), mod=1.0, degree=0, prefiltered=False):
    rs = fb.setup_camera(W, H, common.K_for(W, H), np.eye(4, dtype=np.float32))
    return rs._replace(bg=torch.tensor(bg, dtype=torch.float32, device="cuda"), scale_modifier=mod, sh_degree=degree,
                       prefiltered=prefiltered)


@pytest.mark.parametrize("degree", [0, 1, 2, 3])
def test_sh_colours_forward(cuda, oracle, degree):
    import torch
    W = H = 64
    P, M = 4000, 16
    sc = common.random_view_scene(P, 40 + degree, W, H)
    rng = np.random.default_rng(degree)
    shs = torch.from_numpy(rng.normal(0, 0.5, size=(P, M, 3)).astype(np.float32)).cuda()
    rs = _settings(cuda, torch, W, H, degree=degree)
    t = _t(torch, sc)
    got = _fwd(cuda._C, torch, rs, t, shs=shs, degree=degree)
    ref = _ref_ext()
    if ref is not None:
        # Integers only: in this build the reference extension's SH forward returns an image that depends on what
        # ran before it (all zeros, or not, for the same call), so its colours cannot serve as
        # the expected value.  No reference caller passes SH (colors_precomp always, gaussian.py:1536-1543); the
        # colours are checked against the CPU restatement of forward.cu:20-71 below.
        want = _fwd(ref, torch, rs, t, shs=shs, degree=degree)
        assert got[0] == want[0]
        assert torch.equal(got[2], want[2])
    cam = oracle.make_camera(W, H, rs.tanfovx, rs.tanfovy, rs.viewmatrix.cpu().numpy(), rs.projmatrix.cpu().numpy(),
                             (0, 0, 0), 1.0, degree, M, rs.campos.cpu().numpy(), False)
    o = oracle.Forward(cam, sc["means3D"], sc["opacities"], shs=shs.cpu().numpy(), scales=sc["scales"],
                       rotations=sc["rotations"])
    assert got[0] == o.num_rendered
    img = got[1].cpu().numpy()
    assert (np.abs(img - o.color) > 1e-4 * np.abs(o.color) + 2e-5).mean() < 2e-3
    # degree-0 SH backward: zero dL_dsh (the reference accumulates it only for D > 0, backward.cu:1117)
    if degree == 0:
        dpix = torch.full((3, H, W), 1e-3, device="cuda")
        g = _bwd(cuda._C, torch, rs, t, got, dpix, 1, shs=shs, degree=0)
        assert not g[5].any()
    else:
        # SH backward, degree 1-3 (per-pair fallback, fb_backward.cuh: k_bwd_pairs).
        # (1) power 1 is the true gradient: dL_dsh is checked by a directional finite difference of the forward
        #     (colours are linear in the coefficients wherever they are not clamped).
        g0 = torch.Generator(device="cuda").manual_seed(degree)
        dpix = torch.randn((3, H, W), generator=g0, device="cuda") * 1e-2
        g = _bwd(cuda._C, torch, rs, t, got, dpix, 1, shs=shs, degree=degree)
        nsh = (degree + 1) ** 2
        assert not g[5][:, nsh:].any() and g[5][:, :nsh].abs().sum() > 0
        D = torch.randn(shs.shape, generator=g0, device="cuda")
        D[:, nsh:] = 0
        eps = 1e-3
        lp = (_fwd(cuda._C, torch, rs, t, shs=shs + eps * D, degree=degree)[1].double() * dpix.double()).sum()
        lm = (_fwd(cuda._C, torch, rs, t, shs=shs - eps * D, degree=degree)[1].double() * dpix.double()).sum()
        fd = float((lp - lm) / (2 * eps))
        an = float((g[5].double() * D.double()).sum())
        assert abs(fd - an) <= 2e-2 * abs(an) + 1e-6, (fd, an)
        # (2) against the UNMODIFIED reference extension, including its SH-pointer quirk in dL_dmeans3D,
        #     whenever its own SH forward reproduces the image (it does not always, see above)
        if ref is not None and torch.allclose(want[1], got[1], rtol=1e-4, atol=1e-5):
            for power in (1, 2):
                ga = _bwd(cuda._C, torch, rs, t, got, dpix, power, shs=shs, degree=degree)
                gb = _bwd(ref, torch, rs, t, want, dpix, power, shs=shs, degree=degree)
                for i, (a, b) in enumerate(zip(ga, gb)):
                    a, b = a.double().reshape(P, -1), b.double().reshape(P, -1)
                    scale = b.abs().max()
                    bad = (a - b).abs() > 2e-3 * b.abs() + 2e-6 * scale
                    assert bad.double().mean() < 4e-3, (degree, power, i, int(bad.sum()), bad.numel())


@pytest.mark.parametrize("power", [0, 3, 4, -1])
def test_other_integer_powers_match_reference_extension(cuda, power):
    """backward_power outside {1, 2}: powf(partial, power) per (pixel, Gaussian) pair, summed (backward.cu:1093-1137),
    through the per-pair fallback kernel, against the UNMODIFIED reference extension."""
    import torch
    ref = _ref_ext()
    if ref is None:
        pytest.skip("oracle/_ref/ref_dgr not built")
    W, H, P = 96, 64, 5000
    sc = common.random_view_scene(P, 60 + power, W, H)
    t = _t(torch, sc)
    rs = _settings(cuda, torch, W, H, bg=(0.2, 0.1, 0.4))
    dpix = torch.from_numpy(np.random.default_rng(2).normal(0, 1e-1, size=(3, H, W)).astype(np.float32)).cuda()
    fa, fb_ = _fwd(cuda._C, torch, rs, t), _fwd(ref, torch, rs, t)
    ga, gb = _bwd(cuda._C, torch, rs, t, fa, dpix, power), _bwd(ref, torch, rs, t, fb_, dpix, power)
    for i, (a, b) in enumerate(zip(ga, gb)):
        if i == 5:
            continue
        a, b = a.double().reshape(P, -1), b.double().reshape(P, -1)
        if power == -1:
            # 1 / partial: a pair whose partial is ~0 dominates its sum; compare where the reference is finite and sane
            okm = torch.isfinite(b) & torch.isfinite(a) & (b.abs() < 1e12)
            a, b = a[okm], b[okm]
            bad = (a - b).abs() > 5e-2 * b.abs() + 1e-3 * b.abs().median()
            assert bad.double().mean() < 5e-2, (power, i, float(bad.double().mean()))
            continue
        scale = b.abs().max()
        bad = (a - b).abs() > 2e-3 * b.abs() + 2e-6 * scale
        assert bad.double().mean() < 4e-3, (power, i, int(bad.sum()), bad.numel())


def test_cov3d_precomp_background_and_scale_modifier(cuda):
    import torch
    ref = _ref_ext()
    if ref is None:
        pytest.skip("oracle/_ref/ref_dgr not built")
    W, H, P = 96, 64, 6000
    sc = common.random_view_scene(P, 50, W, H)
    t = _t(torch, sc)
    dpix = torch.from_numpy(np.random.default_rng(1).normal(0, 1e-3, size=(3, H, W)).astype(np.float32)).cuda()
    # (a) scales+rotations with bg != 0 and scale_modifier != 1, both powers
    rs = _settings(cuda, torch, W, H, bg=(0.3, 0.6, 0.1), mod=0.7)
    for power in (1, 2):
        fa, fb_ = _fwd(cuda._C, torch, rs, t), _fwd(ref, torch, rs, t)
        assert fa[0] == fb_[0] and torch.equal(fa[2], fb_[2])
        np.testing.assert_allclose(fa[1].cpu().numpy(), fb_[1].cpu().numpy(), rtol=1e-4, atol=1e-6)
        ga, gb = _bwd(cuda._C, torch, rs, t, fa, dpix, power), _bwd(ref, torch, rs, t, fb_, dpix, power)
        for i, (a, b) in enumerate(zip(ga, gb)):
            if i == 5:
                continue
            a, b = a.double().reshape(P, -1), b.double().reshape(P, -1)
            scale = b.abs().max()
            bad = (a - b).abs() > 2e-3 * b.abs() + 2e-6 * scale
            # the reference's float atomics land in a different order every run: the outlier fraction of this signed,
            # heavily cancelling sum wanders between ~0.5e-3 and ~2e-3 (one suite run in ~12 crossed 2e-3)
            assert bad.double().mean() < 4e-3, (power, i, int(bad.sum()), bad.numel())
    # (b) cov3D_precomp
    rs = _settings(cuda, torch, W, H)
    cov = torch.from_numpy(_cov3d_np(sc["scales"], sc["rotations"])).cuda()
    fa, fb_ = _fwd(cuda._C, torch, rs, t, cov3D=cov), _fwd(ref, torch, rs, t, cov3D=cov)
    assert fa[0] == fb_[0] and torch.equal(fa[2], fb_[2])
    np.testing.assert_allclose(fa[1].cpu().numpy(), fb_[1].cpu().numpy(), rtol=1e-4, atol=1e-6)
    dp = torch.full((3, H, W), 1e-3, device="cuda")
    ga, gb = _bwd(cuda._C, torch, rs, t, fa, dp, 1, cov3D=cov), _bwd(ref, torch, rs, t, fb_, dp, 1, cov3D=cov)
    for i in (0, 1, 2, 3, 4):                       # means2D, colors, opacity, means3D, cov3D
        a, b = ga[i].double().reshape(P, -1), gb[i].double().reshape(P, -1)
        # 3e-3 of the tensor's scale: the worst entry is dL/dz of a Gaussian with a 895-px radius, a signed sum over
        # thousands of pairs that cancels to 1 % of its terms; the reference's float atomics and our moment sums round
        # it differently (block-synchronous and warp-autonomous kernels land at 1.8e-3 and 2.2e-3 of the scale)
        # (run to run the reference alone moves that entry by ~1e-3 of the scale: bound 5e-3)
        assert (a - b).abs().max() <= 5e-3 * b.abs().max() + 1e-12, (i, float((a - b).abs().max() / b.abs().max()))


def test_prefiltered_and_argument_checks(cuda):
    import torch
    W = H = 32
    sc = common.random_view_scene(500, 60, W, H)          # some points are behind the camera
    t = _t(torch, sc)
    rs = _settings(cuda, torch, W, H, prefiltered=True)
    with pytest.raises(RuntimeError):                      # reference: device __trap (auxiliary.h:156-160)
        _fwd(cuda._C, torch, rs, t)
    rs = _settings(cuda, torch, W, H)
    rast = cuda.GaussianRasterizer(raster_settings=rs)
    m2d = torch.zeros_like(t["means3D"])
    shs = torch.zeros((500, 1, 3), device="cuda")
    with pytest.raises(Exception):                         # __init__.py:174-175
        rast(means3D=t["means3D"], means2D=m2d, opacities=t["opacities"], shs=shs, colors_precomp=t["rgb_colors"],
             scales=t["scales"], rotations=t["rotations"])
    with pytest.raises(Exception):
        rast(means3D=t["means3D"], means2D=m2d, opacities=t["opacities"], scales=t["scales"], rotations=t["rotations"])
    with pytest.raises(Exception):                         # __init__.py:177-178
        rast(means3D=t["means3D"], means2D=m2d, opacities=t["opacities"], colors_precomp=t["rgb_colors"],
             scales=t["scales"], rotations=t["rotations"], cov3D_precomp=torch.zeros((500, 6), device="cuda"))
    with pytest.raises(Exception):
        rast(means3D=t["means3D"], means2D=m2d, opacities=t["opacities"], colors_precomp=t["rgb_colors"])


def test_scoring_with_background(cuda):
    """bg != 0 exercises the T_final * bg term of the single-sweep scoring kernel (C' = C + T_final bg)."""
    import torch
    from fisher_nerf_customized_b200 import synth
    W = H = 64
    P, K = 3000, 3
    sc = common.small_scene(P, seed=9)
    rs = cuda.setup_camera(W, H, common.K_for(W, H), np.eye(4, dtype=np.float32))
    rs = rs._replace(bg=torch.tensor([0.2, 0.5, 0.9], device="cuda"))
    t = _t(torch, sc)
    w2c = torch.from_numpy(synth.w2c_from_c2w(synth.make_poses(K, seed=3, box=(6.0, 3.0, 6.0)))).cuda()
    gmap = cuda.GaussianMap(t["means3D"], t["rgb_colors"], t["rotations"], t["opacities"], t["scales"])
    scorer = cuda.FisherScorer(rs, C_H=4, views_per_chunk=2)
    w = torch.rand((P, 4), device="cuda") + 0.1
    got = scorer.score_candidates(gmap, w2c, w).cpu().numpy()
    # the reference caller's loop through the drop-in operator (backward_power = 2)
    want = []
    pts4 = torch.cat((t["means3D"], torch.ones(P, 1, device="cuda")), dim=1)
    for k in range(K):
        tp = (w2c[k] @ pts4.T).T[:, :3].contiguous().requires_grad_(True)
        op = t["opacities"].clone().requires_grad_(True)
        im, _, _ = cuda.GaussianRasterizer(raster_settings=rs, backward_power=2)(
            means3D=tp, means2D=torch.zeros_like(tp, requires_grad=True), opacities=op,
            colors_precomp=t["rgb_colors"], scales=t["scales"], rotations=t["rotations"])
        im.backward(gradient=torch.ones_like(im) * 1e-3)
        Hk = torch.cat([tp.grad, op.grad.reshape(-1, 1)], dim=1)
        want.append(float((Hk * w).sum()))
    np.testing.assert_allclose(got, np.array(want), rtol=1e-3)
