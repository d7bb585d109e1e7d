"""GPU parity of the POp-GS (Hutchinson-probe) variants of the scoring path, SURVEY 8(f) N3:
estimate_diag_JtJ_simple / pose_eval_popgs (models/SLAM/gaussian_object.py:1545-1658, 2046-2109) with the T-/D-optimality
scores (:1705-1718).  Expected values: the reference caller's own arithmetic (one forward, K backward passes with
retain_graph, squares averaged) run through the UNMODIFIED reference extension where it travelled (oracle/_ref),
and through the drop-in autograd operator otherwise.  rtol 1e-3 on the diagonal (atol scaled to the column), 1e-3 on scores.
"""
import numpy as np
import pytest

from fisher_nerf_customized_b200 import synth
from tests import common

pytestmark = pytest.mark.gpu

RTOL = 1e-3


def _ref_ext():
    from oracle import build_ref
    return build_ref.import_ref("ref_dgr")


def _gmap(fb, torch, sc, spatial_order=True):
    t = {k: torch.from_numpy(v).cuda() for k, v in sc.items()}
    return fb.GaussianMap(t["means3D"], t["rgb_colors"], t["rotations"], t["opacities"], t["scales"],
                          spatial_order=spatial_order)


def _caller_diag(C, torch, rs, sc, w2c, probes):
    """The caller arithmetic of estimate_diag_JtJ_simple over a pybind-style `_C` module (reference's or ours)."""
    t = {k: torch.from_numpy(v).cuda() for k, v in sc.items()}
    pts = t["means3D"]
    pts4 = torch.cat([pts, torch.ones(pts.shape[0], 1, device="cuda")], dim=1)
    rel = torch.from_numpy(np.asarray(w2c, dtype=np.float32)).cuda()
    tp = (rel @ pts4.T).T[:, :3].contiguous()
    e = torch.Tensor([])
    fw = C.rasterize_gaussians(rs.bg, tp, t["rgb_colors"], t["opacities"], t["scales"], t["rotations"], rs.scale_modifier,
                               e, rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, rs.image_height, rs.image_width,
                               e, rs.sh_degree, rs.campos, rs.prefiltered)
    R, color, radii, geom, binning, img, depth = fw
    acc = None
    for z in probes:
        g2d, gcol, gop, gmean, gcov, gsh, gsc, grot = C.rasterize_gaussians_backward(
            rs.bg, tp, radii, t["rgb_colors"], t["scales"], t["rotations"], rs.scale_modifier, e, rs.viewmatrix,
            rs.projmatrix, rs.tanfovx, rs.tanfovy, z, e, rs.sh_degree, rs.campos, geom, R, binning, img, 2)
        g = torch.cat([gmean.reshape(-1), gop.reshape(-1), grot.reshape(-1), gsc.reshape(-1)])
        acc = g * g if acc is None else acc + g * g
    return acc / float(len(probes)), int((radii > 0).sum().item())


def _close(got, want, P):
    """Per-column comparison: the four blocks of the concatenated diagonal have very different magnitudes."""
    got, want = got.double().cpu().numpy(), want.double().cpu().numpy()
    bounds = [0, 3 * P, 4 * P, 8 * P, 11 * P]
    for a, b in zip(bounds[:-1], bounds[1:]):
        w, g = want[a:b], got[a:b]
        scale = np.abs(w).max() + 1e-300
        bad = np.abs(g - w) > RTOL * np.abs(w) + RTOL * 1e-3 * scale
        assert bad.mean() < 2e-3, f"block [{a},{b}): {bad.sum()} of {bad.size} entries off"
        assert ((g == 0) != (w == 0)).mean() < 1e-3      # squares of tiny gradients may underflow on one side only


@pytest.mark.parametrize("spatial_order", [True, False])
def test_hutchinson_diag_matches_reference_caller(cuda, spatial_order):
    import torch
    W = H = 64
    P, K = 3000, 3
    sc = common.small_scene(P, seed=11)
    rs = cuda.setup_camera(W, H, common.K_for(W, H), np.eye(4, dtype=np.float32))
    w2c = synth.w2c_from_c2w(synth.make_poses(1, seed=5, box=(6.0, 3.0, 6.0)))[0]
    g = torch.Generator(device="cuda").manual_seed(17)
    probes = [torch.randn((3, H, W), generator=g, device="cuda") for _ in range(K)]
    gmap = _gmap(cuda, torch, sc, spatial_order)
    scorer = cuda.FisherScorer(rs, C_H=11)
    got, vis = scorer.estimate_diag_JtJ(gmap, torch.from_numpy(w2c).cuda(), K=K, probes=probes)
    assert got.shape == (11 * P,)
    ref = _ref_ext()
    want, vis_want = _caller_diag(ref if ref is not None else cuda._C, torch, rs, sc, w2c, probes)
    assert vis == vis_want
    _close(got, want, P)


def test_autograd_retain_graph_probes(cuda):
    """The reference's loop as written (:2077-2103): one forward, K x im.backward(z, retain_graph=True) through the
    drop-in autograd operator, grads zeroed between probes -- equals estimate_diag_JtJ on the same probes."""
    import torch
    W = H = 64
    P, K = 2000, 2
    sc = common.small_scene(P, seed=12)
    rs = cuda.setup_camera(W, H, common.K_for(W, H), np.eye(4, dtype=np.float32))
    t = {k: torch.from_numpy(v).cuda() for k, v in sc.items()}
    g = torch.Generator(device="cuda").manual_seed(3)
    probes = [torch.randn((3, H, W), generator=g, device="cuda") for _ in range(K)]
    tp = t["means3D"].clone().requires_grad_(True)
    rv = dict(means3D=tp, opacities=t["opacities"].requires_grad_(True), rotations=t["rotations"].requires_grad_(True),
              scales=t["scales"].requires_grad_(True), colors_precomp=t["rgb_colors"],
              means2D=torch.zeros_like(tp, requires_grad=True))
    im, radius, _ = cuda.GaussianRasterizer(raster_settings=rs, backward_power=2)(**rv)
    acc = None
    for k, z in enumerate(probes):
        for v in rv.values():
            if v.requires_grad and v.grad is not None:
                v.grad.zero_()
        im.backward(gradient=z, retain_graph=(k < K - 1))
        gg = torch.cat([rv[n].grad.detach().reshape(-1) for n in ("means3D", "opacities", "rotations", "scales")])
        acc = gg * gg if acc is None else acc + gg * gg
    want = acc / float(K)
    scorer = cuda.FisherScorer(rs, C_H=11)
    got, vis = scorer.estimate_diag_JtJ(_gmap(cuda, torch, sc), torch.eye(4).cuda(), K=K, probes=probes)
    assert vis == int((radius > 0).sum())
    _close(got, want, P)


@pytest.mark.parametrize("criterion", ["topt", "dopt"])
def test_pose_eval_popgs_scores(cuda, criterion):
    import torch
    W = H = 64
    P, K, NKF, NC = 2500, 2, 2, 4
    sc = common.small_scene(P, seed=13)
    rs = cuda.setup_camera(W, H, common.K_for(W, H), np.eye(4, dtype=np.float32))
    poses_c2w = synth.make_poses(NKF + NC, seed=9, box=(6.0, 3.0, 6.0))
    w2c = synth.w2c_from_c2w(poses_c2w)
    gmap = _gmap(cuda, torch, sc)
    scorer = cuda.FisherScorer(rs, C_H=11)
    gen = torch.Generator(device="cuda").manual_seed(101)
    scores, poses = scorer.pose_eval_popgs(gmap, torch.from_numpy(poses_c2w[NKF:]).cuda(), torch.from_numpy(w2c[:NKF]),
                                           criterion=criterion, K=K, lam=1e-6, generator=gen)
    assert scores.shape == (NC,) and poses.shape == (NC, 4, 4)
    # the same probes, through the reference extension (or the drop-in _C), with the reference's score formulas
    gen = torch.Generator(device="cuda").manual_seed(101)
    ref = _ref_ext()
    C = ref if ref is not None else cuda._C
    draw = lambda: [torch.randn((3, H, W), generator=gen, device="cuda") for _ in range(K)]
    H_train = None
    for i in range(NKF):
        d, _ = _caller_diag(C, torch, rs, sc, w2c[i], draw())
        H_train = d if H_train is None else H_train + d
    want = []
    for i in range(NKF, NKF + NC):
        w = torch.linalg.inv(torch.from_numpy(poses_c2w[i]).cuda().float()).cpu().numpy()
        d, _ = _caller_diag(C, torch, rs, sc, w, draw())
        Hm = H_train + 1e-6
        if criterion == "topt":
            want.append(-torch.sum(1.0 / torch.clamp(Hm + d, min=1e-12)))
        else:
            want.append(torch.sum(torch.log(torch.clamp(Hm + d, min=1e-12))) - torch.sum(torch.log(torch.clamp(Hm, min=1e-12))))
    want = torch.stack(want).cpu().numpy()
    np.testing.assert_allclose(scores.numpy(), want, rtol=RTOL)
    assert int(np.argmax(scores.numpy())) == int(np.argmax(want))
    with pytest.raises(ValueError):
        scorer.pose_eval_popgs(gmap, torch.from_numpy(poses_c2w[NKF:]).cuda(), torch.from_numpy(w2c[:NKF]), criterion="aopt")
