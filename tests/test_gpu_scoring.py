"""GPU parity of the batched candidate-scoring path (fb_score_candidates / fb_accumulate_hessian)
against the CPU oracle's restatement of pose_eval / compute_Hessian (models/SLAM/gaussian.py:1338-1375,
1503-1570), against the reference's own caller loop run through the drop-in operator, and -- at
BASELINE.json's full size -- through size-independent properties.  rtol 1e-3 on Hessians and scores,
same argmax candidate.
"""
import numpy as np
import pytest

from fisher_nerf_customized_b200 import synth
from tests import common

pytestmark = pytest.mark.gpu

RTOL = 1e-3


def _gmap(fb, torch, sc):
    t = {k: torch.from_numpy(v).cuda() for k, v in sc.items()}
    return fb.GaussianMap(t["means3D"], t["rgb_colors"], t["rotations"], t["opacities"], t["scales"])


def _caller_loop_H(fb, torch, rs, sc, w2c, C_H):
    """compute_Hessian exactly as the reference caller writes it (gaussian.py:1503-1570), through the
    drop-in GaussianRasterizer with backward_power=2."""
    t = {k: torch.from_numpy(v).cuda() for k, v in sc.items()}
    pts = t["means3D"]
    pts4 = torch.cat((pts, torch.ones(pts.shape[0], 1, device="cuda")), dim=1)
    rel = torch.from_numpy(np.asarray(w2c, dtype=np.float32)).cuda()
    tp = (rel @ pts4.T).T[:, :3].contiguous().requires_grad_(True)
    rv = dict(means3D=tp, colors_precomp=t["rgb_colors"].requires_grad_(True), rotations=t["rotations"].requires_grad_(True),
              opacities=t["opacities"].requires_grad_(True), scales=t["scales"].requires_grad_(True),
              means2D=torch.zeros_like(tp, requires_grad=True))
    im, radius, _ = fb.GaussianRasterizer(raster_settings=rs, backward_power=2)(**rv)
    im.backward(gradient=torch.ones_like(im) * 1e-3)
    cols = [tp.grad, rv["opacities"].grad.reshape(-1, 1)]
    if C_H == 11:
        cols += [rv["scales"].grad, rv["rotations"].grad]
    return torch.cat(cols, dim=1)


@pytest.mark.parametrize("C_H", [4, 11])
def test_scores_match_oracle(cuda, oracle, C_H):
    import torch
    W = H = 64
    P, K, NKF = 4000, 6, 3
    sc = common.small_scene(P, seed=3)
    Kmat = common.K_for(W, H)
    rs = cuda.setup_camera(W, H, Kmat, np.eye(4, dtype=np.float32))
    rs_np = oracle.setup_camera_np(W, H, Kmat, np.eye(4))
    poses = synth.w2c_from_c2w(synth.make_poses(K + NKF, seed=7, box=(6.0, 3.0, 6.0)))
    kf, cand = poses[:NKF], poses[NKF:]
    gmap = _gmap(cuda, torch, sc)
    scorer = cuda.FisherScorer(rs, C_H=C_H, views_per_chunk=4)
    # H_train
    H_got = scorer.compute_H_train(gmap, torch.from_numpy(kf).cuda()).cpu().numpy().astype(np.float64)
    H_want = oracle.compute_H_train(rs_np, sc, kf, C_H)
    scale = np.abs(H_want).max(axis=0, keepdims=True)
    bad = np.abs(H_got - H_want) > RTOL * np.abs(H_want) + RTOL * 1e-3 * scale
    assert bad.mean() < 2e-3, f"H_train: {bad.sum()} of {bad.size} entries off"
    np.testing.assert_array_equal(H_got == 0, H_want == 0) if False else None
    # scores with the reference's lambda (gaussian.py:1357) and the yaml's (1e-6)
    for lam in (0.1, 1e-6):
        Hinv = 1.0 / (H_want + lam)
        s_want = oracle.pose_eval(rs_np, sc, cand, Hinv, C_H)
        s_got = scorer.score_candidates(gmap, torch.from_numpy(cand).cuda(),
                                        torch.from_numpy(Hinv.astype(np.float32)).cuda()).cpu().numpy()
        np.testing.assert_allclose(s_got, s_want, rtol=RTOL)
        assert int(np.argmax(s_got)) == int(np.argmax(s_want))


@pytest.mark.parametrize("C_H", [4, 11])
def test_batched_equals_reference_caller_loop(cuda, C_H):
    """One launch sequence for K poses == the reference's K-iteration Python loop over the drop-in
    operator (cur_H = compute_Hessian(w2c); score = sum(cur_H * H_train_inv))."""
    import torch
    W = H = 256
    P, K = 30000, 8
    sc = synth.activate(synth.make_scene(P, seed=5))
    rs = cuda.setup_camera(W, H, common.K_for(W, H), np.eye(4, dtype=np.float32))
    poses = synth.w2c_from_c2w(synth.make_poses(K + 2, seed=8))
    gmap = _gmap(cuda, torch, sc)
    scorer = cuda.FisherScorer(rs, C_H=C_H, views_per_chunk=4)
    H_train = scorer.compute_H_train(gmap, torch.from_numpy(poses[:2]).cuda())
    H_loop = sum(_caller_loop_H(cuda, torch, rs, sc, w, C_H) for w in poses[:2])
    scale = H_loop.abs().amax(dim=0, keepdim=True)
    bad = (H_train - H_loop).abs() > RTOL * H_loop.abs() + RTOL * 1e-3 * scale
    assert bad.float().mean().item() < 1e-3
    Hinv = torch.reciprocal(H_loop + 0.1)
    s_loop = torch.stack([(_caller_loop_H(cuda, torch, rs, sc, w, C_H) * Hinv).sum() for w in poses[2:]]).cpu().numpy()
    s_got = scorer.score_candidates(gmap, torch.from_numpy(poses[2:]).cuda(), Hinv).cpu().numpy()
    np.testing.assert_allclose(s_got, s_loop, rtol=RTOL)
    assert int(np.argmax(s_got)) == int(np.argmax(s_loop))


def test_against_reference_extension_caller_loop(cuda):
    """Same candidates through the UNMODIFIED reference extension driven as pose_eval drives it."""
    import torch
    from oracle import build_ref
    ref_C = build_ref.import_ref("ref_dgr")
    if ref_C is None:
        pytest.skip("oracle/_ref/ref_dgr not built")
    W = H = 256
    P, K = 50000, 6
    sc = synth.activate(synth.make_scene(P, seed=6))
    rs = cuda.setup_camera(W, H, common.K_for(W, H), np.eye(4, dtype=np.float32))
    poses = synth.w2c_from_c2w(synth.make_poses(K + 2, seed=9))
    t = {k: torch.from_numpy(v).cuda() for k, v in sc.items()}
    e = torch.Tensor([])

    def ref_H(w2c):
        pts4 = torch.cat((t["means3D"], torch.ones(P, 1, device="cuda")), dim=1)
        tp = (torch.from_numpy(w2c).cuda() @ pts4.T).T[:, :3].contiguous()
        R, color, radii, geom, binning, img, depth = ref_C.rasterize_gaussians(
            rs.bg, tp, t["rgb_colors"], t["opacities"], t["scales"], t["rotations"], 1.0, e, rs.viewmatrix,
            rs.projmatrix, rs.tanfovx, rs.tanfovy, H, W, e, 0, rs.campos, False)
        g = ref_C.rasterize_gaussians_backward(rs.bg, tp, radii, t["rgb_colors"], t["scales"], t["rotations"], 1.0, e,
                                               rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy,
                                               torch.full((3, H, W), 1e-3, device="cuda"), e, 0, rs.campos, geom, R,
                                               binning, img, 2)
        return torch.cat([g[3], g[2].reshape(-1, 1), g[6], g[7]], dim=1)   # means3D, opacity, scales, rotations

    gmap = _gmap(cuda, torch, sc)
    for C_H in (4, 11):
        cols = slice(0, 4) if C_H == 4 else slice(0, 11)
        H_train = sum(ref_H(w)[:, cols] for w in poses[:2])
        Hinv = torch.reciprocal(H_train + 0.1).contiguous()
        s_ref = torch.stack([(ref_H(w)[:, cols] * Hinv).sum() for w in poses[2:]]).cpu().numpy()
        scorer = cuda.FisherScorer(rs, C_H=C_H, views_per_chunk=8)
        s_got = scorer.score_candidates(gmap, torch.from_numpy(poses[2:]).cuda(), Hinv).cpu().numpy()
        np.testing.assert_allclose(s_got, s_ref, rtol=RTOL)
        assert int(np.argmax(s_got)) == int(np.argmax(s_ref))
        H_got = scorer.compute_H_train(gmap, torch.from_numpy(poses[:2]).cuda())
        scale = H_train.abs().amax(dim=0, keepdim=True)
        bad = (H_got - H_train).abs() > RTOL * H_train.abs() + RTOL * 1e-3 * scale
        assert bad.float().mean().item() < 1e-3


def test_full_size_properties(cuda):
    """BASELINE cfg 3 size (500k Gaussians, 256^2): properties that need no oracle.
    (1) chunking invariance: scores do not depend on views_per_chunk;
    (2) linearity in the weight: score(w1 + 2*w2) == score(w1) + 2*score(w2);
    (3) score(weight = e_c) summed over columns == score(weight = 1) == sum of the Hessian diagonal
        accumulated by fb_accumulate_hessian for the same pose;
    (4) capacity overflow is detected and recovered from (grow and re-run) with identical results."""
    import torch
    W = H = 256
    P, K = 500_000, 12
    sc = synth.activate(synth.make_scene(P))
    rs = cuda.setup_camera(W, H, common.K_for(W, H), np.eye(4, dtype=np.float32))
    w2c = torch.from_numpy(synth.w2c_from_c2w(synth.make_poses(K))).cuda()
    gmap = _gmap(cuda, torch, sc)
    g = torch.Generator(device="cuda").manual_seed(0)
    w1 = torch.rand((P, 4), device="cuda", generator=g)
    w2 = torch.rand((P, 4), device="cuda", generator=g)
    a = cuda.FisherScorer(rs, C_H=4, views_per_chunk=4)
    b = cuda.FisherScorer(rs, C_H=4, views_per_chunk=12, inst_capacity=4096)   # forces the overflow path
    s1 = a.score_candidates(gmap, w2c, w1)
    s1b = b.score_candidates(gmap, w2c, w1)
    assert b.inst_capacity > 4096
    torch.testing.assert_close(s1, s1b, rtol=1e-5, atol=0)
    s2 = a.score_candidates(gmap, w2c, w2)
    s12 = a.score_candidates(gmap, w2c, w1 + 2 * w2)
    torch.testing.assert_close(s12, s1 + 2 * s2, rtol=1e-4, atol=0)
    ones = torch.ones((P, 4), device="cuda")
    s_all = a.score_candidates(gmap, w2c[:2], ones)
    for k in range(2):
        Hk = a.compute_Hessian(gmap, w2c[k])
        torch.testing.assert_close(s_all[k], Hk.double().sum().float(), rtol=1e-3, atol=0)
    assert torch.isfinite(s1).all() and (s1 > 0).all()


def test_spatial_order_is_a_layout_not_a_result(cuda):
    """The Morton-ordered replica (GaussianMap(spatial_order=True), the default) must give the scores and the
    per-Gaussian H_train (in the CALLER's order) of the replica kept in the caller's order."""
    import torch
    W = H = 128
    P, K, NKF = 30000, 8, 2
    sc = common.small_scene(P, seed=11)
    rs = cuda.setup_camera(W, H, common.K_for(W, H), np.eye(4, dtype=np.float32))
    poses = torch.from_numpy(synth.w2c_from_c2w(synth.make_poses(K + NKF, seed=5, box=(6.0, 3.0, 6.0)))).cuda()
    t = {k: torch.from_numpy(v).cuda() for k, v in sc.items()}
    args = (t["means3D"], t["rgb_colors"], t["rotations"], t["opacities"], t["scales"])
    g_sorted, g_plain = cuda.GaussianMap(*args), cuda.GaussianMap(*args, spatial_order=False)
    assert g_sorted.perm is not None and g_plain.perm is None
    assert sorted(g_sorted.perm.tolist()) == list(range(P))
    scorer = cuda.FisherScorer(rs, C_H=4, views_per_chunk=4)
    H_a = scorer.compute_H_train(g_sorted, poses[:NKF])
    H_b = scorer.compute_H_train(g_plain, poses[:NKF])
    scale = H_b.abs().amax(dim=0, keepdim=True)
    assert bool(((H_a - H_b).abs() <= 1e-4 * H_b.abs() + 1e-6 * scale).float().mean() > 0.999)
    H_inv = torch.reciprocal(H_b + 0.1)
    s_a = scorer.score_candidates(g_sorted, poses[NKF:], H_inv).cpu().numpy()
    s_b = scorer.score_candidates(g_plain, poses[NKF:], H_inv).cpu().numpy()
    np.testing.assert_allclose(s_a, s_b, rtol=1e-4)
    assert int(np.argmax(s_a)) == int(np.argmax(s_b))


def test_path_point_eig_equals_step_by_step_loop(cuda):
    """the batched path evaluation equals the reference's sequential loop
    (tester_gaussians_navigation.py:1919-1946) run one pose at a time through compute_Hessian."""
    import torch
    W = H = 64
    P, L, acc = 5000, 13, 5
    sc = common.small_scene(P, seed=21)
    rs = cuda.setup_camera(W, H, common.K_for(W, H), np.eye(4, dtype=np.float32))
    poses = torch.from_numpy(synth.w2c_from_c2w(synth.make_poses(L + 2, seed=8, box=(6.0, 3.0, 6.0)))).cuda()
    gmap = _gmap(cuda, torch, sc)
    scorer = cuda.FisherScorer(rs, C_H=4, views_per_chunk=4)
    H_train = scorer.compute_H_train(gmap, poses[:2])
    path = poses[2:]
    got, total, H_end = scorer.path_point_eig(gmap, path, H_train, 0.1, acc)
    # the reference's loop
    H_path = H_train.clone()
    want, tot = [], 0.0
    for i in range(1, L + 1):
        cur_H = scorer.compute_Hessian(gmap, path[i - 1])
        s = float(torch.sum(cur_H * torch.reciprocal(H_path + 0.1)))
        e = float(np.log(s)) if s > 0 else 0.0
        want.append(e)
        if (i + 1) % acc == 0:
            tot += e
            H_path = H_path + cur_H
    np.testing.assert_allclose(got.cpu().numpy(), np.array(want), rtol=1e-3, atol=1e-4)
    assert abs(float(total) - tot) <= 1e-3 * abs(tot) + 1e-4
    scale = H_path.abs().amax(dim=0, keepdim=True)
    assert bool(((H_end - H_path).abs() <= 1e-4 * H_path.abs() + 1e-6 * scale).all())


def test_map_from_reference_checkpoint(cuda, tmp_path):
    """GaussianMap.from_npz (reference params{t}.npz) == GaussianMap from the same activated tensors; isotropic
    checkpoints (one log-scale per Gaussian, gaussian.py:1532-1533) are broadcast to three."""
    import torch
    raw = synth.make_scene(3000, seed=5, box=(6.0, 3.0, 6.0), focal=32.0, mean_depth=2.0)
    path = tmp_path / "params3.npz"
    np.savez(path, **raw, timestep=np.zeros(3000, np.float32))
    g = cuda.GaussianMap.from_npz(str(path))
    act = synth.activate(raw)
    np.testing.assert_array_equal(g.to_caller_order(g.means3D).cpu().numpy(), act["means3D"])
    np.testing.assert_allclose(g.to_caller_order(g.opacities).cpu().numpy(), act["opacities"].reshape(-1), rtol=1e-6)
    np.testing.assert_allclose(g.to_caller_order(g.rotations).cpu().numpy(), act["rotations"], rtol=1e-5, atol=1e-7)
    rs = cuda.setup_camera(64, 64, common.K_for(64, 64), np.eye(4, dtype=np.float32))
    w2c = torch.from_numpy(synth.w2c_from_c2w(synth.make_poses(3, seed=2, box=(6.0, 3.0, 6.0)))).cuda()
    scorer = cuda.FisherScorer(rs, C_H=4, views_per_chunk=4)
    w = torch.ones((3000, 4), device="cuda")
    s_a = scorer.score_candidates(g, w2c, w).cpu().numpy()
    s_b = scorer.score_candidates(_gmap(cuda, torch, act), w2c, w).cpu().numpy()
    np.testing.assert_allclose(s_a, s_b, rtol=1e-4)
    iso = dict(raw); iso["log_scales"] = raw["log_scales"][:, :1]
    gi = cuda.GaussianMap.from_params({k: torch.from_numpy(np.asarray(v, np.float32)).cuda() for k, v in iso.items()})
    assert tuple(gi.scales.shape) == (3000, 3) and bool((gi.scales[:, 0] == gi.scales[:, 2]).all())


def test_batched_edge_cases(cuda):
    """No candidates; an image that is not a multiple of the tile size; a workspace too small for the instance list
    (the driver grows it and re-runs); tile lists of thousands of entries with only 21 distinct depths (the sort's
    equal-key fallback) -- each against the reference caller's loop through the drop-in operator."""
    import torch
    # (1) K = 0
    W, H = 100, 70
    sc = common.random_view_scene(6000, 51, W, H)
    rs = cuda.setup_camera(W, H, common.K_for(W, H), np.eye(4, dtype=np.float32))
    gmap = _gmap(cuda, torch, sc)
    scorer = cuda.FisherScorer(rs, C_H=4, views_per_chunk=4)
    Hinv = torch.ones((6000, 4), device="cuda")
    assert scorer.score_candidates(gmap, torch.zeros((0, 4, 4), device="cuda"), Hinv).shape == (0,)
    assert not scorer.compute_H_train(gmap, torch.zeros((0, 4, 4), device="cuda")).any()
    # (2) ragged image, K not a multiple of the chunk, (3) capacity far too small at first
    eye = np.eye(4, dtype=np.float32)
    poses = np.stack([eye] + [common.look_at_w2c((0.2 * i, 0.1 * i, -0.3 * i), (0.0, 0.0, 2.0)) for i in range(1, 6)]).astype(np.float32)
    want = torch.stack([(_caller_loop_H(cuda, torch, rs, sc, w, 4) * Hinv).sum() for w in poses]).cpu().numpy()
    got = scorer.score_candidates(gmap, torch.from_numpy(poses).cuda(), Hinv).cpu().numpy()
    np.testing.assert_allclose(got, want, rtol=RTOL)
    small = cuda.FisherScorer(rs, C_H=4, views_per_chunk=4, inst_capacity=256)
    got2 = small.score_candidates(gmap, torch.from_numpy(poses).cuda(), Hinv).cpu().numpy()
    assert small.inst_capacity > 256
    np.testing.assert_array_equal(got2, got)
    # (4) depth ties and long lists; caller's order kept so that ties break as in the reference
    W = H = 32
    P = 9000
    sc = common.random_view_scene(P, 31, W, H)
    rng = np.random.default_rng(31)
    sc["means3D"][:, :2] = rng.uniform(-0.3, 0.3, (P, 2)).astype(np.float32)
    sc["means3D"][:, 2] = np.round(rng.uniform(1.0, 3.0, P), 1).astype(np.float32)
    sc["scales"][:] = 0.02
    rs = cuda.setup_camera(W, H, common.K_for(W, H), eye)
    t = {k: torch.from_numpy(v).cuda() for k, v in sc.items()}
    gmap = cuda.GaussianMap(t["means3D"], t["rgb_colors"], t["rotations"], t["opacities"], t["scales"], spatial_order=False)
    for C_H in (4, 11):
        scorer = cuda.FisherScorer(rs, C_H=C_H, views_per_chunk=2)
        H_want = _caller_loop_H(cuda, torch, rs, sc, eye, C_H)
        H_got = scorer.compute_Hessian(gmap, torch.from_numpy(eye).cuda())
        scale = H_want.abs().amax(dim=0, keepdim=True)
        bad = (H_got - H_want).abs() > RTOL * H_want.abs() + RTOL * 1e-3 * scale
        assert bad.float().mean().item() < 2e-3
        Hinv = torch.reciprocal(H_want + 0.1)
        s = scorer.score_candidates(gmap, torch.from_numpy(np.stack([eye, eye])).cuda(), Hinv).cpu().numpy()
        np.testing.assert_allclose(s, np.full(2, float((H_want * Hinv).sum())), rtol=RTOL)


@pytest.mark.parametrize("C_H", [4, 11])
def test_chunk_lanes_do_not_change_scores(cuda, C_H):
    """Two chunk lanes (consecutive chunks on two streams, two workspace halves) vs one: the same bits, the same
    status counters; also with an odd number of chunks, a ragged last chunk, and a capacity overflow + regrow."""
    import torch
    W = H = 64
    P, K = 6000, 23
    sc = common.small_scene(P, seed=31)
    rs = cuda.setup_camera(W, H, common.K_for(W, H), np.eye(4, dtype=np.float32))
    w2c = torch.from_numpy(synth.w2c_from_c2w(synth.make_poses(K, seed=32, box=(6.0, 3.0, 6.0)))).cuda()
    gmap = _gmap(cuda, torch, sc)
    g = torch.Generator(device="cuda").manual_seed(1)
    H_inv = torch.rand((P, C_H), generator=g, device="cuda") + 0.1
    out = {}
    for lanes in (1, 2):
        scorer = cuda.FisherScorer(rs, C_H=C_H, views_per_chunk=4, chunk_lanes=lanes, inst_capacity=2000)  # forces a regrow
        s = scorer.score_candidates(gmap, w2c, H_inv)
        # a second call on a side stream: the lanes fork from and join the CALLER's stream
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            s2 = scorer.score_candidates(gmap, w2c, H_inv)
        side.synchronize()
        out[lanes] = (s.cpu().numpy(), s2.cpu().numpy(), scorer.last_instances, scorer.last_visible)
    assert np.array_equal(out[1][0], out[2][0]) and np.array_equal(out[2][0], out[2][1])
    assert out[1][2:] == out[2][2:]
    assert np.all(out[2][0] > 0)


def test_plan_best_path_vs_reference_extension_loop(cuda):
    """all candidate paths of a planning step in one launch sequence per accumulation level, against the
    reference's loop (tester_gaussians_navigation.py:1664-1721) run pose by pose through the UNMODIFIED reference
    extension: per-path totals within rtol 1e-3 (they are sums of logs), the same best path."""
    import torch
    from oracle import build_ref
    ref_C = build_ref.import_ref("ref_dgr")
    if ref_C is None:
        pytest.skip("oracle/_ref/ref_dgr not built")
    W = H = 128
    P, acc, lam = 40000, 5, 0.1
    sc = synth.activate(synth.make_scene(P, seed=13, focal=64.0))
    rs = cuda.setup_camera(W, H, common.K_for(W, H), np.eye(4, dtype=np.float32))
    t = {k: torch.from_numpy(v).cuda() for k, v in sc.items()}
    gmap = _gmap(cuda, torch, sc)
    lengths = [14, 9, 3, 11, 19, 5]                      # 3, 2, 0, 2, 4, 1 accumulation levels (steps 4, 9, 14, 19)
    all_poses = synth.w2c_from_c2w(synth.make_poses(sum(lengths) + 3, seed=17))
    kf, rest = all_poses[:3], all_poses[3:]
    paths, o = [], 0
    for L in lengths:
        paths.append(rest[o:o + L]); o += L
    final_EIGs = [0.3, 1.7, 0.9, 0.2, 1.1, 0.5]
    e = torch.Tensor([])
    pts4 = torch.cat((t["means3D"], torch.ones(P, 1, device="cuda")), dim=1)
    dpix = torch.full((3, H, W), 1e-3, device="cuda")

    def ref_H(w2c):
        tp = (torch.from_numpy(w2c).cuda() @ pts4.T).T[:, :3].contiguous()
        R, color, radii, geom, binning, img, depth = ref_C.rasterize_gaussians(
            rs.bg, tp, t["rgb_colors"], t["opacities"], t["scales"], t["rotations"], 1.0, e, rs.viewmatrix,
            rs.projmatrix, rs.tanfovx, rs.tanfovy, H, W, e, 0, rs.campos, False)
        g = ref_C.rasterize_gaussians_backward(rs.bg, tp, radii, t["rgb_colors"], t["scales"], t["rotations"], 1.0, e,
                                               rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, dpix, e, 0,
                                               rs.campos, geom, R, binning, img, 2)
        return torch.cat([g[3], g[2].reshape(-1, 1)], dim=1), torch.eye(6, device="cuda")

    H_train = sum(ref_H(w)[0] for w in kf)
    pose_w, point_w, end_w = 0.2, 1.0, 1.0
    want = []
    for path, fe in zip(paths, final_EIGs):              # the reference's loop, statement by statement
        H_path = H_train.clone()
        total, curr_action = 0, []
        for w2c in path:
            cur_H, pose_H = ref_H(w2c)
            H_inv_path = torch.reciprocal(H_path + lam)
            point_EIG = torch.log(torch.sum(cur_H * H_inv_path))
            pose_EIG = torch.log(torch.linalg.det(pose_H))
            curr_action.append(1)
            total += pose_w * pose_EIG.item()
            if (len(curr_action) + 1) % acc == 0:
                total += point_w * point_EIG.item()
                H_path = H_path + cur_H
        want.append(total / len(curr_action) + end_w * fe)
    scorer = cuda.FisherScorer(rs, C_H=4, views_per_chunk=8)
    got, best, point = scorer.plan_best_path(gmap, paths, final_EIGs, H_train, lam, acc, pose_w, point_w, end_w)
    np.testing.assert_allclose(np.array(got), np.array(want), rtol=1e-3, atol=1e-5)
    assert best == int(np.argmax(want))
    assert int(torch.isfinite(point).sum()) == sum(sum(1 for i in range(1, L + 1) if (i + 1) % acc == 0) for L in lengths) == 12
    # the object variant drops the point term (:1929, :1936) and only the end term decides
    got_o, best_o, _ = scorer.plan_best_path(gmap, paths, final_EIGs, H_train, lam, acc, pose_w, point_w, end_w,
                                             object_variant=True)
    np.testing.assert_allclose(np.array(got_o), np.array(final_EIGs), rtol=1e-6)
    assert best_o == 1


def test_morton_order_and_incremental_map(cuda):
    """The Morton permutation (one code kernel + one sort) groups neighbours; a map grown by `append_rows` (activated
    tail rows, lazy re-sort) scores exactly like a map rebuilt from the full parameter set; a ResidentParams store keeps
    an attached map in step."""
    import torch
    from fisher_nerf_customized_b200.scoring import morton_order
    pts = torch.from_numpy(synth.make_points(20000, seed=3)).cuda()
    perm = morton_order(pts)
    assert sorted(perm.tolist()) == list(range(20000))
    s = pts[perm]
    d_sorted = (s[1:] - s[:-1]).norm(dim=1).median()
    d_plain = (pts[1:] - pts[:-1]).norm(dim=1).median()
    assert float(d_sorted) < 0.1 * float(d_plain)
    assert morton_order(pts[:1]).tolist() == [0]
    assert sorted(morton_order(torch.zeros((5, 3), device="cuda")).tolist()) == [0, 1, 2, 3, 4]
    flat = pts.clone(); flat[:, 1] = 0.25
    assert sorted(morton_order(flat).tolist()) == list(range(20000))
    # incremental growth == rebuild
    W = H = 64
    P0, steps, n_new = 6000, 5, 900
    raw = synth.make_scene(P0 + steps * n_new, seed=41, box=(6.0, 3.0, 6.0), focal=32.0, mean_depth=2.0)
    params = {k: torch.from_numpy(np.asarray(v, np.float32)).cuda() for k, v in raw.items()}
    head = {k: v[:P0] for k, v in params.items()}
    rs = cuda.setup_camera(W, H, common.K_for(W, H), np.eye(4, dtype=np.float32))
    poses = torch.from_numpy(synth.w2c_from_c2w(synth.make_poses(5, seed=6, box=(6.0, 3.0, 6.0)))).cuda()
    grown = cuda.GaussianMap.from_params(head, capacity=P0 + 2 * n_new)        # capacity forces one regrow too
    grown.resort_fraction = 0.2                                                # and one lazy re-sort on the way
    for i in range(steps):
        r0 = P0 + i * n_new
        grown.append_rows({k: v[:r0 + n_new] for k, v in params.items()}, r0, n_new)
    full = cuda.GaussianMap.from_params(params)
    assert grown.P == full.P == P0 + steps * n_new and sorted(grown.perm.tolist()) == list(range(grown.P))
    scorer = cuda.FisherScorer(rs, C_H=4, views_per_chunk=4)
    H_a, H_b = scorer.compute_H_train(grown, poses[:2]), scorer.compute_H_train(full, poses[:2])
    scale = H_b.abs().amax(dim=0, keepdim=True)
    assert bool(((H_a - H_b).abs() <= 1e-4 * H_b.abs() + 1e-6 * scale).float().mean() > 0.999)
    w = torch.reciprocal(H_b + 0.1)
    np.testing.assert_allclose(scorer.score_candidates(grown, poses[2:], w).cpu().numpy(),
                               scorer.score_candidates(full, poses[2:], w).cpu().numpy(), rtol=1e-4)
    with pytest.raises(RuntimeError):
        grown.append_rows(params, 5, 1)                                        # rows must start behind the live ones


def test_heavy_tile_lists_split_sort_and_segmented_sweep(cuda):
    """Tile lists of >= 16384 keys take their own paths: the sort partitions them by depth and sorts the pieces with
    other CTAs, the pair sweep cuts them into segments chained through global memory.  120k small, faint Gaussians on a
    32 x 32 image (4 tiles, ~200 distinct depths so that the pieces are full of ties) put > 30k keys on every tile with
    pixels that stay live over several segments and others that stop inside one.  Checked against the caller's loop
    through the drop-in rasteriser, whose kernels (single-view binning, k_render_warp / k_bwd_warp) share none of this."""
    import torch
    W = H = 32
    P = 120_000
    rng = np.random.default_rng(77)
    sc = common.random_view_scene(P, 77, W, H)
    z = np.round(rng.uniform(1.0, 3.0, P), 2).astype(np.float32)
    sc["means3D"][:, 0] = rng.uniform(-0.8, 0.8, P).astype(np.float32)
    sc["means3D"][:, 1] = rng.uniform(-0.8, 0.8, P).astype(np.float32)
    sc["means3D"][:, 2] = z
    sc["scales"][:] = rng.uniform(0.01, 0.03, (P, 3)).astype(np.float32)
    sc["opacities"][:] = rng.uniform(0.01, 0.08, (P, 1)).astype(np.float32)
    eye = np.eye(4, dtype=np.float32)
    rs = cuda.setup_camera(W, H, common.K_for(W, H), eye)
    t = {k: torch.from_numpy(v).cuda() for k, v in sc.items()}
    gmap = cuda.GaussianMap(t["means3D"], t["rgb_colors"], t["rotations"], t["opacities"], t["scales"], spatial_order=False)
    H_want = _caller_loop_H(cuda, torch, rs, sc, eye, 4)
    scorer = cuda.FisherScorer(rs, C_H=4, views_per_chunk=2)
    Hinv = torch.reciprocal(H_want + 0.1)
    poses = torch.from_numpy(np.stack([eye, eye, eye])).cuda()
    s = scorer.score_candidates(gmap, poses, Hinv).cpu().numpy()
    assert scorer.last_status[2] >= 16384, scorer.last_status          # the longest tile list: the heavy paths ran
    np.testing.assert_allclose(s, np.full(3, float((H_want * Hinv).sum())), rtol=RTOL)
    assert s[0] == s[1] == s[2]                                         # same view, same bits: the combine order is fixed
    # run to run as well
    s2 = scorer.score_candidates(gmap, poses, Hinv).cpu().numpy()
    np.testing.assert_array_equal(s, s2)
    # an instance capacity that is far too small at first: the chunk is flagged, nothing of it runs, the driver grows and repeats
    tight = cuda.FisherScorer(rs, C_H=4, views_per_chunk=2, inst_capacity=1000)
    s3 = tight.score_candidates(gmap, poses, Hinv).cpu().numpy()
    assert tight.inst_capacity > 1000
    np.testing.assert_array_equal(s3, s)
    # the H_train path sorts the same lists through the split sort
    H_got = scorer.compute_Hessian(gmap, torch.from_numpy(eye).cuda())
    scale = H_want.abs().amax(dim=0, keepdim=True)
    bad = (H_got - H_want).abs() > RTOL * H_want.abs() + RTOL * 1e-3 * scale
    assert bad.float().mean().item() < 2e-3
