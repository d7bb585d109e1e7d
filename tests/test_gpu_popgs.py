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


# ---- block form (estimate_block_JtJ / t_opt_blocks / d_opt_blocks / pose_eval_popgs_blocks) ----------------------

def _caller_blocks(C, torch, rs, sc, w2c, probes, use_opacity=True, use_rot=True, use_scale=True):
    """estimate_block_JtJ's arithmetic (gaussian_object.py:2147-2174) over a pybind-style `_C` module."""
    t = {k: torch.from_numpy(v).cuda() for k, v in sc.items()}
    pts = t["means3D"]
    pts4 = torch.cat([pts, torch.ones(pts.shape[0], 1, device="cuda")], dim=1)
    rel = torch.from_numpy(np.asarray(w2c, dtype=np.float32)).cuda()
    tp = (rel @ pts4.T).T[:, :3].contiguous()
    e = torch.Tensor([])
    R, color, radii, geom, binning, img, depth = C.rasterize_gaussians(
        rs.bg, tp, t["rgb_colors"], t["opacities"], t["scales"], t["rotations"], rs.scale_modifier, e, rs.viewmatrix,
        rs.projmatrix, rs.tanfovx, rs.tanfovy, rs.image_height, rs.image_width, e, rs.sh_degree, rs.campos,
        rs.prefiltered)
    vis_idx = torch.where(radii > 0)[0]
    H_blocks = None
    for z in probes:
        g2d, gcol, gop, gmean, gcov, gsh, gsc, grot = C.rasterize_gaussians_backward(
            rs.bg, tp, radii, t["rgb_colors"], t["scales"], t["rotations"], rs.scale_modifier, e, rs.viewmatrix,
            rs.projmatrix, rs.tanfovx, rs.tanfovy, z, e, rs.sh_degree, rs.campos, geom, R, binning, img, 2)
        cols = [gmean] + ([gop] if use_opacity else []) + ([grot] if use_rot else []) + ([gsc] if use_scale else [])
        G = torch.cat([c.reshape(c.shape[0], -1) for c in cols], dim=1)
        Gv = G[vis_idx.long(), :]
        o = Gv.unsqueeze(2) * Gv.unsqueeze(1)
        H_blocks = o if H_blocks is None else H_blocks + o
    return H_blocks / float(len(probes)), vis_idx.long()


@pytest.mark.parametrize("flags", [(True, True, True), (False, True, False), (True, False, True), (False, False, False)])
def test_block_outer_accumulate_is_the_reference_expression(cuda, flags):
    """fb_block_outer_accumulate vs `H_blocks += Gv.unsqueeze(2) * Gv.unsqueeze(1)` ... `/ float(K)` on the same
    gradient tensors: bit-identical (same rounding points)."""
    import torch
    from fisher_nerf_customized_b200 import _lib
    use_op, use_rot, use_sc = flags
    L = _lib.lib()
    g = torch.Generator(device="cuda").manual_seed(5)
    P, K = 5000, 3
    rows = torch.nonzero(torch.rand(P, generator=g, device="cuda") < 0.4).reshape(-1)
    Nv = rows.numel()
    d = 3 + use_op + 4 * use_rot + 3 * use_sc
    got = torch.zeros((Nv, d, d), device="cuda")
    want = torch.zeros((Nv, d, d), device="cuda")
    for k in range(K):
        gm, go = torch.randn(P, 3, generator=g, device="cuda"), torch.randn(P, 1, generator=g, device="cuda")
        gr, gs = torch.randn(P, 4, generator=g, device="cuda"), torch.randn(P, 3, generator=g, device="cuda")
        cols = [gm] + ([go] if use_op else []) + ([gr] if use_rot else []) + ([gs] if use_sc else [])
        Gv = torch.cat(cols, dim=1)[rows]
        want += Gv.unsqueeze(2) * Gv.unsqueeze(1)
        rc = L.fb_block_outer_accumulate(Nv, rows.data_ptr(), gm.data_ptr(), go.data_ptr() if use_op else None,
                                         gr.data_ptr() if use_rot else None, gs.data_ptr() if use_sc else None,
                                         float(K) if k == K - 1 else 0.0, got.data_ptr(), None)
        assert rc == 0
    want = want / float(K)
    torch.cuda.synchronize()
    assert torch.equal(got, want)


@pytest.mark.parametrize("d", [3, 4, 7, 8, 10, 11])
@pytest.mark.parametrize("criterion", ["topt", "dopt"])
def test_block_opt_matches_torch_linalg(cuda, d, criterion):
    """t_opt_blocks / d_opt_blocks (gaussian_object.py:1720-1732): the FP32 matrices the reference forms, factorised in
    FP64 here -> rtol 1e-9 against torch.linalg in FP64 on those matrices, and rtol 1e-3 against the reference's own
    FP32 expression where FP32 can resolve it (lam = 1e-2)."""
    import torch
    g = torch.Generator(device="cuda").manual_seed(100 + d)
    n = 1237                                                  # not a multiple of the CTA size
    def blocks(rank):
        v = torch.randn(n, rank, d, generator=g, device="cuda")
        return torch.einsum("nri,nrj->nij", v, v).contiguous()
    Hm, J = blocks(4), blocks(2)                              # rank-deficient, like a few probes x a few keyframes
    rs = cuda.setup_camera(32, 32, common.K_for(32, 32), np.eye(4, dtype=np.float32))
    scorer = cuda.FisherScorer(rs, C_H=11)
    I32 = torch.eye(d, device="cuda")
    for lam in (1e-2, 1e-6):
        if criterion == "topt":
            got, terms = scorer._block_opt(Hm, J, lam, 0, per_block=True)
            M32 = Hm + J + lam * I32
            want64 = -torch.einsum("bii->b", torch.linalg.inv(M32.double()))
            want32 = -torch.einsum("bii->", torch.linalg.inv(M32))
            assert torch.allclose(scorer.t_opt_blocks(Hm, J, lam), got)
        else:
            got, terms = scorer._block_opt(Hm, J, lam, 1, per_block=True)
            M0 = Hm + lam * I32
            M1 = M0 + J
            want64 = torch.linalg.slogdet(M1.double())[1] - torch.linalg.slogdet(M0.double())[1]
            want32 = (torch.linalg.slogdet(M1)[1] - torch.linalg.slogdet(M0)[1]).sum()
            assert torch.allclose(scorer.d_opt_blocks(Hm, J, lam), got)
        np.testing.assert_allclose(terms.double().cpu().numpy(), want64.cpu().numpy(), rtol=1e-6, atol=1e-6)
        np.testing.assert_allclose(float(got), float(want64.sum()), rtol=1e-6)
        if lam == 1e-2:
            np.testing.assert_allclose(float(got), float(want32), rtol=RTOL)


def test_block_opt_gathers_and_edge_cases(cuda):
    import torch
    rs = cuda.setup_camera(32, 32, common.K_for(32, 32), np.eye(4, dtype=np.float32))
    scorer = cuda.FisherScorer(rs, C_H=11)
    g = torch.Generator(device="cuda").manual_seed(1)
    d = 11
    v = torch.randn(50, 3, d, generator=g, device="cuda")
    Hm = torch.einsum("nri,nrj->nij", v, v).contiguous()
    w = torch.randn(70, 2, d, generator=g, device="cuda")
    J = torch.einsum("nri,nrj->nij", w, w).contiguous()
    im = torch.tensor([0, 7, 49, 3], device="cuda")
    ij = torch.tensor([69, 0, 5, 5], device="cuda")
    got = scorer._block_opt(Hm, J, 1e-3, 0, im, ij)
    want = scorer.t_opt_blocks(Hm[im], J[ij], 1e-3)
    assert float(got) == float(want)
    # empty set of blocks: the sum over nothing
    assert float(scorer.t_opt_blocks(Hm[:0], J[:0], 1e-3)) == 0.0
    # intersect_sorted == np.intersect1d(..., return_indices=True)
    from fisher_nerf_customized_b200.scoring import intersect_sorted
    a = torch.tensor([1, 4, 5, 9, 12, 40], device="cuda")
    b = torch.tensor([0, 4, 9, 10, 40, 41, 50], device="cuda")
    ia, ib = intersect_sorted(a, b)
    _, wa, wb = np.intersect1d(a.cpu().numpy(), b.cpu().numpy(), return_indices=True)
    assert ia.cpu().tolist() == wa.tolist() and ib.cpu().tolist() == wb.tolist()
    ia, ib = intersect_sorted(a, torch.tensor([100, 200], device="cuda"))
    assert ia.numel() == 0 and ib.numel() == 0


@pytest.mark.parametrize("spatial_order", [True, False])
def test_estimate_block_JtJ_matches_reference_caller(cuda, spatial_order):
    import torch
    W = H = 64
    P, K = 3000, 3
    sc = common.small_scene(P, seed=21)
    rs = cuda.setup_camera(W, H, common.K_for(W, H), np.eye(4, dtype=np.float32))
    w2c = synth.w2c_from_c2w(synth.make_poses(1, seed=6, box=(6.0, 3.0, 6.0)))[0]
    g = torch.Generator(device="cuda").manual_seed(23)
    probes = [torch.randn((3, H, W), generator=g, device="cuda") for _ in range(K)]
    scorer = cuda.FisherScorer(rs, C_H=11)
    got, vis = scorer.estimate_block_JtJ(_gmap(cuda, torch, sc, spatial_order), torch.from_numpy(w2c).cuda(), K=K,
                                         probes=probes)
    ref = _ref_ext()
    want, vis_want = _caller_blocks(ref if ref is not None else cuda._C, torch, rs, sc, w2c, probes)
    assert torch.equal(vis, vis_want)                       # same visible set, caller's indices, ascending
    assert got.shape == want.shape == (vis.numel(), 11, 11)
    got, want = got.double().cpu().numpy(), want.double().cpu().numpy()
    assert np.array_equal(got, got.transpose(0, 2, 1))     # symmetric by construction
    # entry (r, c) scales like |g_r||g_c|: compare against the geometric mean of the two diagonal entries
    diag = np.sqrt(np.abs(np.einsum("bii->bi", want)))
    scale = diag[:, :, None] * diag[:, None, :] + 1e-300
    colmax = np.abs(want).reshape(-1, 121).max(axis=0).reshape(11, 11)
    bad = np.abs(got - want) > RTOL * np.abs(want) + 2 * RTOL * scale + RTOL * 1e-3 * colmax
    assert bad.mean() < 2e-3, f"{bad.sum()} of {bad.size} block entries off"
    # a sub-selection of the columns is the matching sub-block
    got2, vis2 = scorer.estimate_block_JtJ(_gmap(cuda, torch, sc, spatial_order), torch.from_numpy(w2c).cuda(), K=K,
                                           probes=probes, use_rot=False, use_scale=True, use_opacity=False)
    keep = [0, 1, 2, 8, 9, 10]
    assert torch.equal(vis2, vis)
    # (not bit-equal: the two backward runs add their atomics in different orders)
    sub, sub_scale = got[:, keep][:, :, keep], scale[:, keep][:, :, keep]
    bad = np.abs(got2.double().cpu().numpy() - sub) > RTOL * np.abs(sub) + 2 * RTOL * sub_scale
    assert bad.mean() < 2e-3


@pytest.mark.parametrize("criterion", ["topt", "dopt"])
def test_pose_eval_popgs_blocks_scores(cuda, criterion):
    """pose_eval_popgs_blocks end to end vs the reference's loop (compute_H_train_blocks' positional truncation,
    np.intersect1d pairing, torch.linalg on the FP32 matrices) over the reference extension, same probes.
    lam = 1.0: the block entries of the two extensions agree to 1e-3 of the block's scale (atomics add in different
    orders, in the reference too), which is an ABSOLUTE perturbation of a block's small eigenvalues, and 1/(mu + lam)
    amplifies it by scale/lam; with lam above the block scale the scores agree to rtol 1e-3.  At the reference's
    default lam = 1e-6 the criterion is that noise amplified 1e4-fold in EITHER implementation, so the small-lam
    case is tested on identical blocks instead (test_score_blocks_on_identical_blocks, test_block_opt_*)."""
    import torch
    W = H = 64
    P, K, NKF, NC = 2500, 2, 2, 3
    lam = 1.0
    sc = common.small_scene(P, seed=13)
    rs = cuda.setup_camera(W, H, common.K_for(W, H), np.eye(4, dtype=np.float32))
    poses_c2w = synth.make_poses(NKF + NC, seed=9, box=(6.0, 3.0, 6.0))
    w2c = synth.w2c_from_c2w(poses_c2w)
    scorer = cuda.FisherScorer(rs, C_H=11)
    gen = torch.Generator(device="cuda").manual_seed(77)
    scores, poses = scorer.pose_eval_popgs_blocks(_gmap(cuda, torch, sc), torch.from_numpy(poses_c2w[NKF:]).cuda(),
                                                  torch.from_numpy(w2c[:NKF]), criterion=criterion, K=K, lam=lam,
                                                  generator=gen)
    assert scores.shape == (NC,) and poses.shape == (NC, 4, 4)
    gen = torch.Generator(device="cuda").manual_seed(77)
    ref = _ref_ext()
    C = ref if ref is not None else cuda._C
    draw = lambda: [torch.randn((3, H, W), generator=gen, device="cuda") for _ in range(K)]
    per_kf = [_caller_blocks(C, torch, rs, sc, w2c[i], draw()) for i in range(NKF)]
    Hm, vis_ref = _reference_H_train_blocks(per_kf)
    want = []
    for i in range(NKF, NKF + NC):
        wc = torch.linalg.inv(torch.from_numpy(poses_c2w[i]).cuda().float()).cpu().numpy()
        Jb, cv = _caller_blocks(C, torch, rs, sc, wc, draw())
        want.append(_reference_block_score(torch, Hm, vis_ref, Jb, cv, criterion, lam))
    np.testing.assert_allclose(scores.numpy(), np.array(want), rtol=RTOL)
    assert int(np.argmax(scores.numpy())) == int(np.argmax(want))


def _reference_H_train_blocks(per_kf):
    Hm, vis_ref = None, None
    for Hb, vi in per_kf:                                    # compute_H_train_blocks, :1574-1583
        if Hm is None:
            Hm, vis_ref = Hb, vi
        else:
            n = min(Hm.shape[0], Hb.shape[0])
            Hm, vis_ref = Hm[:n] + Hb[:n], vis_ref[:n]
    return Hm, vis_ref


def _reference_block_score(torch, Hm, vis_ref, Jb, cv, criterion, lam, dtype=None):
    """The loop body of pose_eval_popgs_blocks (:1671-1698) with t_opt_blocks / d_opt_blocks (:1720-1732)."""
    _, it, ic = np.intersect1d(vis_ref.cpu().numpy(), cv.cpu().numpy(), return_indices=True)
    if len(it) == 0:
        return float("-inf")
    Hb, J = Hm[torch.from_numpy(it).cuda()], Jb[torch.from_numpy(ic).cuda()]
    I = torch.eye(Hm.shape[-1], device="cuda")
    dt = dtype or torch.float64
    if criterion == "topt":
        return float(-torch.einsum("bii->", torch.linalg.inv((Hb + J + lam * I).to(dt))))
    M0 = Hb + lam * I
    return float((torch.linalg.slogdet((M0 + J).to(dt))[1] - torch.linalg.slogdet(M0.to(dt))[1]).sum())


@pytest.mark.parametrize("criterion", ["topt", "dopt"])
def test_score_blocks_on_identical_blocks(cuda, criterion):
    """Pairing + truncation + scoring at the reference's small lam on ONE set of rendered blocks: our
    compute_H_train_blocks rule and score_blocks vs the reference's loop body on the same tensors."""
    import torch
    W = H = 64
    P, K = 2500, 2
    sc = common.small_scene(P, seed=14)
    rs = cuda.setup_camera(W, H, common.K_for(W, H), np.eye(4, dtype=np.float32))
    w2c = synth.w2c_from_c2w(synth.make_poses(5, seed=10, box=(6.0, 3.0, 6.0)))
    scorer = cuda.FisherScorer(rs, C_H=11)
    gmap = _gmap(cuda, torch, sc)
    gen = torch.Generator(device="cuda").manual_seed(5)
    blocks = [scorer.estimate_block_JtJ(gmap, torch.from_numpy(w).cuda(), K=K, generator=gen) for w in w2c]
    Hm, vis_ref = _reference_H_train_blocks(blocks[:2])
    for lam in (1e-6, 1e-3):
        for Jb, cv in blocks[2:]:
            got = float(scorer.score_blocks(Hm, vis_ref, Jb, cv, criterion, lam))
            want = _reference_block_score(torch, Hm, vis_ref, Jb, cv, criterion, lam)
            np.testing.assert_allclose(got, want, rtol=1e-5)
    # a candidate that shares no Gaussian with the prior
    far = torch.tensor([P + 5, P + 9], device="cuda")
    assert float(scorer.score_blocks(Hm, vis_ref, blocks[2][0][:2], far, criterion, 1e-6)) == float("-inf")
