#!/usr/bin/env python
"""tests/bench_configs.py -- BASELINE.json's five configs on ONE B200, each next to the reference where one exists.
(Lives under tests/ because it runs the checkers in oracle/ beside the product.)

  python tests/bench_configs.py [--only 1,2,4,5] [--loop-steps 200]

Prints one JSON line per config.  bench.py is config 3 (the headline metric) and is not repeated here.
Every timed GPU region uses CUDA events after warm-up; every line carries the parity check made at that size
(bit-exact integers, tolerances of the north star for floating point).  The reference arms are the reference's
own code: its CUDA extensions rebuilt for sm_100a (oracle/_ref, only if present) and its numpy
count_visible_points (oracle.count_visible_points restates it operation by operation; pinned by tests/golden/vis_a).
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def ev_time(fn, iters, warm=3):
    import torch
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record(); b.synchronize()
    return a.elapsed_time(b) / iters


def cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count()


def cfg1():
    """count_visible_points: 1M points x 64 poses at 256x256 -- GPU batched counter vs the numpy CPU path."""
    import torch
    import fisher_nerf_customized_b200 as fb
    from fisher_nerf_customized_b200 import synth
    from oracle import oracle as O
    N, K = 1_000_000, 64
    pts = synth.make_points(N)
    c2w = synth.make_poses(K).astype(np.float64)
    intr = synth.intrinsics(256, 256).astype(np.float32)
    dpts = torch.from_numpy(pts).cuda()
    ms = ev_time(lambda: fb.count_visible_batch(dpts, c2w, intr, 256), 20)
    t0 = time.perf_counter()
    got = fb.count_visible_batch(torch.from_numpy(pts).cuda(), c2w, intr, 256).cpu().numpy()     # host points in, counts out
    e2e = time.perf_counter() - t0
    t0 = time.perf_counter()
    want = np.array([len(O.count_visible_points(pts, c2w[k], intr, 256)) for k in range(K)])
    cpu = time.perf_counter() - t0
    return {"config": "1: count_visible_points, 1M points x 64 poses, 256x256", "unit": "point-poses/s",
            "b200_resident": N * K / (ms * 1e-3), "b200_ms": ms, "b200_e2e_host_points": N * K / e2e,
            "cpu_numpy_reference": N * K / cpu, "cpu_s": cpu, "cores": cores(),
            "parity": {"counts_equal": bool((got == want).all()), "sum_counts": int(want.sum())}}


def cfg2():
    """Single-view GaussianRasterizer forward + backward, 200k Gaussians at 256x256, vs the reference extension."""
    import torch
    import fisher_nerf_customized_b200 as fb
    from fisher_nerf_customized_b200 import synth
    from oracle import build_ref
    P, W, H = 200_000, 256, 256
    sc = synth.activate(synth.make_scene(P))
    t = {k: torch.from_numpy(v).cuda() for k, v in sc.items()}
    w2c = torch.from_numpy(synth.w2c_from_c2w(synth.make_poses(1))[0]).cuda()
    pts4 = torch.cat((t["means3D"], torch.ones(P, 1, device="cuda")), dim=1)
    tp = (w2c @ pts4.T).T[:, :3].contiguous()
    rs = fb.setup_camera(W, H, synth.intrinsics(W, H), np.eye(4, dtype=np.float32))
    e = torch.Tensor([])
    dpix = torch.ones((3, H, W), device="cuda") * 1e-3
    out = {"config": "2: single-view rasterizer fwd+bwd, 200k Gaussians, 256x256", "unit": "ms per fwd+bwd"}

    def run(C, power):
        R, color, radii, geom, binning, img, depth = C.rasterize_gaussians(
            rs.bg, tp, t["rgb_colors"], t["opacities"], t["scales"], t["rotations"], 1.0, e, rs.viewmatrix, rs.projmatrix,
            rs.tanfovx, rs.tanfovy, H, W, e, 0, rs.campos, False)
        g = C.rasterize_gaussians_backward(rs.bg, tp, radii, t["rgb_colors"], t["scales"], t["rotations"], 1.0, e,
                                           rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, dpix, e, 0, rs.campos,
                                           geom, R, binning, img, power)
        return R, color, radii, g
    for power in (1, 2):
        out[f"b200_power{power}"] = ev_time(lambda: run(fb._C, power), 20)
    ref = build_ref.import_ref("ref_dgr")
    if ref is not None:
        for power in (1, 2):
            out[f"reference_power{power}"] = ev_time(lambda: run(ref, power), 20)
        Ra, ca, ra, ga = run(fb._C, 2)
        Rb, cb, rb, gb = run(ref, 2)
        gm_a, gm_b = ga[3].double(), gb[3].double()          # dL_dmeans3D (power 2: Hessian diagonal columns 0-2)
        scale = gm_b.abs().max()
        out["parity"] = {"num_rendered_equal": Ra == Rb, "radii_equal": bool((ra == rb).all()),
                         "image_max_abs_diff": float((ca - cb).abs().max()),
                         "H_means_frac_outside_rtol1e-3": float(((gm_a - gm_b).abs() > 1e-3 * gm_b.abs() + 1e-6 * scale).double().mean())}
    else:
        out["reference"] = "oracle/_ref not built"
    return out


def cfg4():
    """HM3D-scale map: 3M Gaussians x 1024 candidates, plus distCUDA2 over the 3M means."""
    import torch
    import fisher_nerf_customized_b200 as fb
    from fisher_nerf_customized_b200 import synth
    from oracle import build_ref
    P, K, W, H = 3_000_000, 1024, 256, 256
    box = (30.0, 3.0, 30.0)
    sc = synth.activate(synth.make_scene(P, box=box))
    t = {k: torch.from_numpy(v).cuda() for k, v in sc.items()}
    gmap = fb.GaussianMap(t["means3D"], t["rgb_colors"], t["rotations"], t["opacities"], t["scales"])
    rs = fb.setup_camera(W, H, synth.intrinsics(W, H), np.eye(4, dtype=np.float32))
    scorer = fb.FisherScorer(rs, C_H=4, views_per_chunk=64)
    kf = torch.from_numpy(synth.w2c_from_c2w(synth.make_poses(8, seed=synth.SEED_POSES + 1, box=box))).cuda()
    H_inv = torch.reciprocal(scorer.compute_H_train(gmap, kf) + 0.1)
    w2c = torch.from_numpy(synth.w2c_from_c2w(synth.make_poses(K, box=box))).cuda()
    ms = ev_time(lambda: scorer.score_candidates(gmap, w2c, H_inv), 3, warm=2)
    scores = scorer.score_candidates(gmap, w2c, H_inv)
    out = {"config": "4: 3M Gaussians x 1024 candidates at 256x256 + distCUDA2 over 3M points",
           "scoring_views_per_s": K / (ms * 1e-3), "scoring_ms": ms, "instances_per_view": scorer.last_instances / K,
           "scores_finite_positive": bool(torch.isfinite(scores).all() and (scores >= 0).all()),
           "argmax": int(scores.argmax())}
    # plain layout gives the same scores (layout is not a result)
    g2 = fb.GaussianMap(t["means3D"], t["rgb_colors"], t["rotations"], t["opacities"], t["scales"], spatial_order=False)
    s2 = scorer.score_candidates(g2, w2c[:64], H_inv)
    out["max_rel_diff_vs_unsorted_layout"] = float(((scores[:64] - s2).abs() / s2.abs().clamp_min(1e-30)).max())
    del g2
    knn_ms = ev_time(lambda: fb.distCUDA2(t["means3D"]), 5, warm=2)
    out["distCUDA2_ms"] = knn_ms
    out["distCUDA2_points_per_s"] = P / (knn_ms * 1e-3)
    ref = build_ref.import_ref("ref_knn")
    if ref is not None:
        d_ref = ref.distCUDA2(t["means3D"])
        torch.cuda.synchronize()
        t0 = time.perf_counter(); d_ref = ref.distCUDA2(t["means3D"]); torch.cuda.synchronize()
        out["distCUDA2_reference_ms"] = (time.perf_counter() - t0) * 1e3
        out["distCUDA2_bit_identical_to_reference"] = bool(torch.equal(fb.distCUDA2(t["means3D"]), d_ref))
    return out


def cfg5(steps):
    """Synthetic active-mapping loop: every step appends 3276 Gaussians and counts visible points at the current
    pose; every 10th step refreshes H_train over the keyframes (every 4th frame, the 32 most recent) and scores
    256 candidates."""
    import torch
    import fisher_nerf_customized_b200 as fb
    from fisher_nerf_customized_b200 import synth
    W = H = 256
    add, P0, cap = 3276, 100_000, 3_000_000
    total = min(cap, P0 + steps * add)
    sc = synth.activate(synth.make_scene(total))
    t = {k: torch.from_numpy(v).cuda() for k, v in sc.items()}
    poses = synth.make_poses(steps + 256, seed=77)
    w2c_all = torch.from_numpy(synth.w2c_from_c2w(poses)).cuda()
    rs = fb.setup_camera(W, H, synth.intrinsics(W, H), np.eye(4, dtype=np.float32))
    intr = synth.intrinsics(W, H).astype(np.float32)
    scorer = fb.FisherScorer(rs, C_H=4, views_per_chunk=64)
    cand = w2c_all[steps:steps + 256]
    picks, vis_counts = [], []
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for s in range(steps):
        P = min(total, P0 + s * add)                                     # map update: rows [0, P) are live
        vis_counts.append(fb.count_visible_batch(t["means3D"][:P], poses[s:s + 1].astype(np.float64), intr, W))
        if s % 10 == 9:
            gmap = fb.GaussianMap(t["means3D"][:P], t["rgb_colors"][:P], t["rotations"][:P], t["opacities"][:P], t["scales"][:P])
            kfs = list(range(0, s + 1, 4))[-32:]
            H_inv = torch.reciprocal(scorer.compute_H_train(gmap, w2c_all[kfs]) + 0.1)
            picks.append(scorer.score_candidates(gmap, cand, H_inv).argmax())
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    return {"config": f"5: active-mapping loop, {steps} steps (map {P0} -> {min(total, P0 + (steps - 1) * add)} Gaussians), "
                      "visibility count every step, H_train refresh (<= 32 keyframes) + 256-candidate scoring every 10th",
            "steps_per_s": steps / dt, "wall_s": dt, "scoring_rounds": len(picks),
            "picks_head": [int(p) for p in picks[:8]], "visible_last": int(vis_counts[-1][0])}


def cfg5a():
    """Map-update half of config 5: densify one 256x256 RGB-D frame (5 % of the pixels new, tester_navigator.py:380-381)
    into a 3M-Gaussian map.  Reference arm: get_pointcloud's tensor program + initialize_new_params + the torch.cat
    growth of add_new_gaussians (gaussian.py:83-137, 299-318, 399-401) written out with torch on the same GPU."""
    import torch
    import fisher_nerf_customized_b200 as fb
    from tests.golden.make_golden import frame_inputs
    H = W = 256
    P = 3_000_000
    color, depth, K, w2c, _ = frame_inputs(H, W, 5)
    rng = np.random.default_rng(6)
    mask = (rng.random(H * W) < 0.05) & (depth.reshape(-1) > 0.01)
    tc, td, tK, tw, tm = [torch.from_numpy(np.ascontiguousarray(x)).cuda() for x in (color, depth, K, w2c, mask)]
    store = fb.ResidentParams(capacity=P + 64 * 4000)
    store.n = P

    def ours():
        store.n = P
        store.append_from_frame(tc, td, tK, tw, mask=tm)
    params = {k: torch.zeros((P, c), device="cuda") for k, c in fb.ResidentParams.KEYS}

    def ref():
        CX, CY, FX, FY = tK[0][2], tK[1][2], tK[0][0], tK[1][1]
        xg, yg = torch.meshgrid(torch.arange(0, W).cuda().float(), torch.arange(0, H).cuda().float(), indexing="xy")
        xx, yy = ((xg - CX) / FX).reshape(-1), ((yg - CY) / FY).reshape(-1)
        z = td[0].reshape(-1)
        pts4 = torch.cat((torch.stack((xx * z, yy * z, z), dim=-1), torch.ones(H * W, 1).cuda()), dim=1)
        pts = (torch.inverse(tw).cuda() @ pts4.T).T[:, :3]
        msd = ((1 * z / ((FX + FY) / 2)) ** 2)[tm]
        cld = torch.cat((pts, torch.permute(tc, (1, 2, 0)).reshape(-1, 3)), -1)[tm]
        n = cld.shape[0]
        rots = torch.zeros((n, 4), device="cuda"); rots[:, 0] = 1
        new = {"means3D": cld[:, :3], "rgb_colors": cld[:, 3:6], "unnorm_rotations": rots,
               "logit_opacities": torch.zeros((n, 1), device="cuda"),
               "log_scales": torch.tile(torch.log(torch.sqrt(msd))[..., None], (1, 3))}
        return {k: torch.cat((params[k], v), dim=0) for k, v in new.items()}
    t_ours, t_ref = ev_time(ours, 50), ev_time(ref, 20)
    grown = ref()
    got = store.params()
    ok = all(torch.allclose(got[k][P:], grown[k][P:], rtol=2e-6, atol=2e-6) for k in got)
    return {"config": "5a: map update, densify one 256x256 RGB-D frame into a 3M-Gaussian map", "new_gaussians": int(mask.sum()),
            "b200_ms": t_ours, "reference_torch_program_ms": t_ref, "parity": {"new_rows_allclose_2e-6": bool(ok)}}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", default="1,2,4,5,5a")
    ap.add_argument("--loop-steps", type=int, default=200)
    args = ap.parse_args()
    from fisher_nerf_customized_b200 import _lib
    _lib.lib()
    todo = {"1": cfg1, "2": cfg2, "4": cfg4, "5": lambda: cfg5(args.loop_steps), "5a": cfg5a}
    for k in args.only.split(","):
        print(json.dumps(todo[k.strip()]()), flush=True)


if __name__ == "__main__":
    main()
