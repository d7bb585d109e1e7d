"""GPU parity at the BENCHMARKED sizes against the UNMODIFIED reference extension (oracle/_ref/ref_dgr.so, the
reference's sources rebuilt for sm_100a), driven exactly as pose_eval drives it
(models/SLAM/gaussian.py:1354-1375 -> compute_Hessian :1503-1570):

  cfg 3  500k Gaussians x ALL 256 candidates, H_train from the 32 keyframes, C_H in {4, 11}, lambda in {0.1, 1e-6}
  cfg 4  3M Gaussians, a 32-candidate sample, H_train from 8 keyframes, C_H = 4, lambda = 0.1

Every score is held to rtol 1e-3 and the argmax candidate must be the same.  The reference extension's own
run-to-run noise (float atomics) is measured too, and the tolerance fractions used elsewhere are tied to it.
"""
import numpy as np
import pytest

from fisher_nerf_customized_b200 import synth
from tests import common

pytestmark = pytest.mark.gpu

RTOL = 1e-3
W = H = 256


def _ref(torch):
    from oracle import build_ref
    ref_C = build_ref.import_ref("ref_dgr")
    if ref_C is None:
        pytest.skip("oracle/_ref/ref_dgr not built")
    return ref_C


def _ref_H(torch, ref_C, rs, t, pts4, w2c, dpix):
    """cur_H[P, 11] of one pose through the reference extension: means3D(3), opacity(1), scales(3), rotations(4)
    (gaussian.py:1554-1560; gaussian_object.py:2024-2032)."""
    e = torch.Tensor([])
    tp = (w2c @ pts4.T).T[:, :3].contiguous()
    R, color, radii, geom, binning, img, depth = ref_C.rasterize_gaussians(
        rs.bg, tp, t["rgb_colors"], t["opacities"], t["scales"], t["rotations"], 1.0, e, rs.viewmatrix,
        rs.projmatrix, rs.tanfovx, rs.tanfovy, H, W, e, 0, rs.campos, False)
    g = ref_C.rasterize_gaussians_backward(rs.bg, tp, radii, t["rgb_colors"], t["scales"], t["rotations"], 1.0, e,
                                           rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, dpix, e, 0,
                                           rs.campos, geom, R, binning, img, 2)
    return torch.cat([g[3], g[2].reshape(-1, 1), g[6], g[7]], dim=1), R, radii


def _setup(cuda, torch, P, box):
    sc = synth.activate(synth.make_scene(P, box=box))
    t = {k: torch.from_numpy(v).cuda() for k, v in sc.items()}
    rs = cuda.setup_camera(W, H, common.K_for(W, H), np.eye(4, dtype=np.float32))
    gmap = cuda.GaussianMap(t["means3D"], t["rgb_colors"], t["rotations"], t["opacities"], t["scales"])
    pts4 = torch.cat((t["means3D"], torch.ones(P, 1, device="cuda")), dim=1)
    dpix = torch.full((3, H, W), 1e-3, device="cuda")
    return sc, t, rs, gmap, pts4, dpix


def test_cfg3_all_256_candidates_vs_reference_extension(cuda):
    import torch
    ref_C = _ref(torch)
    P, K, NKF = 500_000, 256, 32
    sc, t, rs, gmap, pts4, dpix = _setup(cuda, torch, P, (10.0, 3.0, 10.0))
    kf = torch.from_numpy(synth.w2c_from_c2w(synth.make_poses(NKF, seed=synth.SEED_POSES + 1))).cuda()
    cand = torch.from_numpy(synth.w2c_from_c2w(synth.make_poses(K, seed=synth.SEED_POSES))).cuda()
    H_train = torch.zeros((P, 11), device="cuda")
    R_kf = 0
    for w in kf:
        h, R, _ = _ref_H(torch, ref_C, rs, t, pts4, w, dpix)
        H_train += h
        R_kf += R
    cases = [(C_H, lam) for C_H in (4, 11) for lam in (0.1, 1e-6)]
    Hinv = {(C_H, lam): torch.reciprocal(H_train[:, :C_H] + lam).contiguous() for C_H, lam in cases}
    s_ref = {c: [] for c in cases}
    R_tot, vis_tot = 0, 0
    for w in cand:
        h, R, radii = _ref_H(torch, ref_C, rs, t, pts4, w, dpix)
        R_tot += R
        vis_tot += int((radii > 0).sum())
        for (C_H, lam) in cases:
            s_ref[(C_H, lam)].append(torch.sum(h[:, :C_H] * Hinv[(C_H, lam)]))
    for C_H in (4, 11):
        scorer = cuda.FisherScorer(rs, C_H=C_H, views_per_chunk=64)
        # H_train of the batched path against the reference's sum of per-view tensors
        H_got = scorer.compute_H_train(gmap, kf)
        ref = H_train[:, :C_H]
        scale = ref.abs().amax(dim=0, keepdim=True)
        bad = (H_got - ref).abs() > RTOL * ref.abs() + RTOL * 1e-3 * scale
        assert bad.float().mean().item() < 1e-3, f"H_train C_H={C_H}: {int(bad.sum())} of {bad.numel()} entries off"
        assert scorer.last_instances == R_kf                       # the reference's num_rendered, summed: bit-exact
        for lam in (0.1, 1e-6):
            want = torch.stack(s_ref[(C_H, lam)]).cpu().numpy()
            got = scorer.score_candidates(gmap, cand, Hinv[(C_H, lam)]).cpu().numpy()
            rel = np.abs(got - want) / np.abs(want)
            assert rel.max() <= RTOL, f"C_H={C_H} lambda={lam}: max rel diff {rel.max():.2e} at candidate {rel.argmax()}"
            assert int(np.argmax(got)) == int(np.argmax(want))
            assert scorer.last_instances == R_tot and scorer.last_visible == vis_tot


def test_cfg4_sample_vs_reference_extension(cuda):
    import torch
    ref_C = _ref(torch)
    P, K, NKF = 3_000_000, 32, 8
    box = (30.0, 3.0, 30.0)
    sc, t, rs, gmap, pts4, dpix = _setup(cuda, torch, P, box)
    kf = torch.from_numpy(synth.w2c_from_c2w(synth.make_poses(NKF, seed=synth.SEED_POSES + 1, box=box))).cuda()
    cand = torch.from_numpy(synth.w2c_from_c2w(synth.make_poses(K, seed=synth.SEED_POSES, box=box))).cuda()
    H_train = sum(_ref_H(torch, ref_C, rs, t, pts4, w, dpix)[0][:, :4] for w in kf)
    Hinv = torch.reciprocal(H_train + 0.1).contiguous()
    want = torch.stack([torch.sum(_ref_H(torch, ref_C, rs, t, pts4, w, dpix)[0][:, :4] * Hinv) for w in cand]).cpu().numpy()
    scorer = cuda.FisherScorer(rs, C_H=4, views_per_chunk=32)
    got = scorer.score_candidates(gmap, cand, Hinv).cpu().numpy()
    rel = np.abs(got - want) / np.abs(want)
    assert rel.max() <= RTOL, f"max rel diff {rel.max():.2e}"
    assert int(np.argmax(got)) == int(np.argmax(want))


def test_visible_capacity_overflow_is_recovered(cuda):
    """Per-view records are sized by the visible capacity; a view that exceeds it raises FB_E_CAPACITY on the
    device, the driver grows the capacity and re-runs: same bits as a run that had room from the start."""
    import torch
    P, K = 200_000, 6
    sc, t, rs, gmap, pts4, dpix = _setup(cuda, torch, P, (10.0, 3.0, 10.0))
    cand = torch.from_numpy(synth.w2c_from_c2w(synth.make_poses(K, seed=3))).cuda()
    w = torch.rand((P, 4), device="cuda", generator=torch.Generator(device="cuda").manual_seed(0)) + 0.1
    roomy = cuda.FisherScorer(rs, C_H=4, views_per_chunk=4, vis_capacity=P)
    tight = cuda.FisherScorer(rs, C_H=4, views_per_chunk=4, vis_capacity=1000)
    a = roomy.score_candidates(gmap, cand, w)
    b = tight.score_candidates(gmap, cand, w)
    assert tight.vis_capacity > 1000 and tight.vis_capacity >= tight.last_status[5]
    assert torch.equal(a, b)
    assert roomy.last_instances == tight.last_instances and roomy.last_visible == tight.last_visible


def test_reference_extension_noise_floor(cuda):
    """The reference accumulates its Hessian with float atomics, so two runs on one input differ.  This measures the
    fraction of H entries (of Gaussians with radius > 0) that move by more than rtol 1e-3 between two runs of the
    UNMODIFIED extension, and the same fraction for this repo's drop-in operator against it: the bound the raster
    tests use (`frac`) must not be looser than twice what the reference does to itself (plus one entry in 10^4 for
    the alpha ~ 1/255 decisions that differ by rounding)."""
    import torch
    ref_C = _ref(torch)
    P = 200_000
    sc, t, rs, gmap, pts4, dpix = _setup(cuda, torch, P, (10.0, 3.0, 10.0))
    w2c = torch.from_numpy(synth.w2c_from_c2w(synth.make_poses(1, seed=11))).cuda()[0]
    a, _, radii = _ref_H(torch, ref_C, rs, t, pts4, w2c, dpix)
    b, _, _ = _ref_H(torch, ref_C, rs, t, pts4, w2c, dpix)
    vis = radii > 0
    scale = a.abs().amax(dim=0, keepdim=True)

    def frac_off(x, y):
        bad = (x - y).abs() > RTOL * y.abs() + RTOL * 1e-3 * scale
        return float(bad[vis].float().mean())

    self_noise = frac_off(a, b)
    e = torch.Tensor([])
    tp = (w2c @ pts4.T).T[:, :3].contiguous()
    R, color, radii2, geom, binning, img, depth = cuda.rasterizer._C.rasterize_gaussians(
        rs.bg, tp, t["rgb_colors"], t["opacities"], t["scales"], t["rotations"], 1.0, e, rs.viewmatrix,
        rs.projmatrix, rs.tanfovx, rs.tanfovy, H, W, e, 0, rs.campos, False)
    g = cuda.rasterizer._C.rasterize_gaussians_backward(
        rs.bg, tp, radii2, t["rgb_colors"], t["scales"], t["rotations"], 1.0, e, rs.viewmatrix, rs.projmatrix,
        rs.tanfovx, rs.tanfovy, dpix, e, 0, rs.campos, geom, R, binning, img, 2)
    ours = torch.cat([g[3], g[2].reshape(-1, 1), g[6], g[7]], dim=1)
    assert torch.equal(radii2, radii)
    vs_ref = frac_off(ours, a)
    print(f"reference self-noise fraction {self_noise:.2e}; this repo vs reference {vs_ref:.2e}")
    assert vs_ref <= 2.0 * self_noise + 1e-4, (self_noise, vs_ref)
