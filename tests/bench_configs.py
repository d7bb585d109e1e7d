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


def cfg3_variants():
    """BASELINE cfg 3's other settings (bench.py is C_H = 4, lambda = 0.1): C_H = 11 and lambda = 1e-6, resident views/s."""
    import torch
    import fisher_nerf_customized_b200 as fb
    from fisher_nerf_customized_b200 import synth
    P, K, W, H = 500_000, 256, 256, 256
    sc = synth.activate(synth.make_scene(P))
    t = {k: torch.from_numpy(v).cuda() for k, v in sc.items()}
    gmap = fb.GaussianMap(t["means3D"], t["rgb_colors"], t["rotations"], t["opacities"], t["scales"])
    rs = fb.setup_camera(W, H, synth.intrinsics(W, H), np.eye(4, dtype=np.float32))
    kf = torch.from_numpy(synth.w2c_from_c2w(synth.make_poses(32, seed=synth.SEED_POSES + 1))).cuda()
    w2c = torch.from_numpy(synth.w2c_from_c2w(synth.make_poses(K))).cuda()
    out = {"config": "3 (variants): 500k Gaussians x 256 candidates at 256x256, resident"}
    for C_H in (4, 11):
        scorer = fb.FisherScorer(rs, C_H=C_H, views_per_chunk=64)
        H_train = scorer.compute_H_train(gmap, kf)
        out[f"C_H{C_H}_H_train_keyframes_per_s"] = 32 / (ev_time(lambda: scorer.compute_H_train(gmap, kf), 5, warm=2) * 1e-3)
        for lam in (0.1, 1e-6):
            H_inv = torch.reciprocal(H_train + lam)
            ms = ev_time(lambda: scorer.score_candidates(gmap, w2c, H_inv), 5, warm=2)
            out[f"C_H{C_H}_lambda{lam}_views_per_s"] = K / (ms * 1e-3)
            out[f"C_H{C_H}_lambda{lam}_argmax"] = int(scorer.score_candidates(gmap, w2c, H_inv).argmax())
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


def _room_frames(poses_c2w, W, H, K, rng, device):
    """Synthetic RGB-D frames of the 10 x 3 x 10 m box room seen from each pose: z-depth by ray / box intersection,
    colour from the hit point (torch, on the device; data generation, outside every timed region)."""
    import torch
    c2w = torch.from_numpy(np.asarray(poses_c2w, np.float32)).to(device)
    ys, xs = torch.meshgrid(torch.arange(H, device=device, dtype=torch.float32),
                            torch.arange(W, device=device, dtype=torch.float32), indexing="ij")
    dc = torch.stack([(xs - K[0, 2]) / K[0, 0], (ys - K[1, 2]) / K[1, 1], torch.ones_like(xs)], -1)      # [H,W,3], unit z
    d = torch.einsum("nij,hwj->nhwi", c2w[:, :3, :3], dc)
    o = c2w[:, None, None, :3, 3]
    half = torch.tensor([5.0, 1.5, 5.0], device=device)
    t = torch.where(d > 0, (half - o) / d, (-half - o) / d)
    t = torch.where(torch.isfinite(t), t, torch.full_like(t, 1e9))
    depth = t.amin(dim=-1)                                                                               # [N,H,W]
    hit = o + depth[..., None] * d
    color = (0.5 + 0.5 * torch.sin(hit * 3.0)).permute(0, 3, 1, 2).contiguous()
    return color, depth[:, None].contiguous()


def cfg5(steps):
    """BASELINE config 5 as specified: a synthetic active-mapping loop.  Every step the map owner (rank 0)
    densifies a synthetic 256x256 RGB-D frame into its parameter store (`ResidentParams.append_from_frame`: 3276 new
    Gaussians = 5 % of the pixels, tester_navigator.py:380-381, until the map holds 3M), the new rows are broadcast to
    the replicas (`broadcast_appended`: the delta only), the attached scoring maps follow (activated rows appended,
    lazy Morton re-sort), and the visible points of the current pose are counted; every 10th step (`map_every`,
    configs/base_config.py:47) H_train is refreshed over the keyframes (every 4th frame, the 32 most recent, dealt
    over the ranks + all-reduce) and 256 candidates are scored (split over the ranks + all-gather): one
    ShardedScorer.pose_eval.  Run under torchrun for N > 1 GPUs."""
    import torch
    import fisher_nerf_customized_b200 as fb
    from fisher_nerf_customized_b200 import synth
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    W = H = 256
    add, P0, cap = 3276, 100_000, 3_000_000
    Kmat = synth.intrinsics(W, H).astype(np.float32)
    poses = synth.make_poses(steps + 256, seed=77)
    w2c_all = torch.from_numpy(synth.w2c_from_c2w(poses)).to(dev)
    c2w_all = torch.from_numpy(poses).to(dev)
    rs = fb.setup_camera(W, H, Kmat, np.eye(4, dtype=np.float32), device=dev)
    scorer = fb.FisherScorer(rs, C_H=4, views_per_chunk=64)
    sharded = fb.ShardedScorer(scorer)
    # the initial map: the same seeded 100k Gaussians on every rank (the full 56 * P broadcast happens once, at init)
    raw = synth.make_scene(P0)
    store = fb.ResidentParams(capacity=cap + 2 * add, device=dev)
    for k in store.buf:
        store.buf[k][:P0] = torch.from_numpy(np.asarray(raw[k], np.float32)).to(dev)
    store.n = P0
    sharded.broadcast_map([store.buf[k][:P0] for k in sorted(store.buf)], src=0)
    gmap = store.attach(fb.GaussianMap.from_params(store.params(), capacity=cap + 2 * add))
    # synthetic frames (owner only needs them) and the per-frame pixel selections, generated before the clock starts
    rng = np.random.default_rng(5)
    nf = min(steps, 64)                                          # 64 distinct frames, cycled
    Kt = torch.from_numpy(Kmat).to(dev)
    color, depth = _room_frames(poses[:nf], W, H, Kmat, rng, dev) if rank == 0 else (None, None)
    masks = torch.zeros((nf, H * W), dtype=torch.bool, device=dev)
    if rank == 0:
        for f in range(nf):
            masks[f, torch.from_numpy(rng.choice(H * W, size=add, replace=False)).to(dev)] = True
    cand = torch.from_numpy(synth.make_poses(256, seed=78)).to(dev)     # the same candidates whatever `steps` is
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    picks, vis_last = [], None
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for s in range(steps):
        f = s % nf
        ev[s][0].record()
        r0 = n = None
        if rank == 0:
            if store.n + add <= cap:
                r0, n = store.append_from_frame(color[f], depth[f], Kt, w2c_all[f], mask=masks[f], time_idx=s)
            else:
                r0, n = store.n, 0
        sharded.broadcast_appended(store, r0, n, src=0)          # delta only; attached maps follow on every rank
        ev[s][1].record()
        vis_last = fb.count_visible_batch(store.buf["means3D"][:store.n], poses[s:s + 1].astype(np.float64), Kmat, W)
        if s % 10 == 9:
            kfs = list(range(0, s + 1, 4))[-32:]
            scores, best = sharded.pose_eval(gmap, cand, w2c_all[kfs], 0.1)
            picks.append(best)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    dt = time.perf_counter() - t0
    upd_ms = sum(a.elapsed_time(b) for a, b in ev) / steps
    # replicas must be bit-identical: wrap-around checksums of the raw parameter bits
    chk = torch.stack([store.buf[k][:store.n].contiguous().view(torch.int32).sum(dtype=torch.int64) for k in sorted(store.buf)]
                      + [torch.tensor(store.n, dtype=torch.int64, device=dev),
                         gmap.means3D.contiguous().view(torch.int32).sum(dtype=torch.int64)])
    same = True
    if world > 1:
        allc = torch.empty((world, chk.numel()), dtype=torch.int64, device=dev)
        dist.all_gather_into_tensor(allc, chk)
        same = bool((allc == allc[0:1]).all())
    out = {"config": f"5: active-mapping loop, {steps} steps on {world} GPU(s): densify a 256x256 RGB-D frame ({add} new Gaussians, "
                     f"map {P0} -> {store.n}), delta broadcast, visibility count every step; every 10th step H_train over <= 32 "
                     "keyframes + 256-candidate scoring (ShardedScorer.pose_eval)",
           "n_gpus": world, "steps_per_s": steps / dt, "wall_s": dt, "map_update_ms_per_step": upd_ms,
           "scoring_rounds": len(picks), "picks_head": [int(p) for p in picks[:8]], "visible_last": int(vis_last[0]),
           "final_gaussians": int(store.n), "replicas_bit_identical": same, "map_resorts": "lazy (tail > 5 %)"}
    if world > 1:
        dist.destroy_process_group()
    return out if rank == 0 else None


def cfg5_reference(steps):
    """The same loop the way the reference runs it, on ONE GPU, for a bounded number of steps: get_pointcloud's tensor
    program + torch.cat growth of every parameter array (gaussian.py:83-137, 299-318, 399-401), numpy
    count_visible_points on the host (tester_navigator.py:147-180), and every 10th step compute_H_train + pose_eval through
    the UNMODIFIED reference extension, one pose at a time (gaussian.py:1338-1375, 1503-1570)."""
    import torch
    import torch.nn.functional as F
    from fisher_nerf_customized_b200 import synth
    from oracle import build_ref, oracle as O
    ref_C = build_ref.import_ref("ref_dgr")
    if ref_C is None:
        return {"config": "5 (reference arm)", "unavailable": "oracle/_ref/ref_dgr.so not built"}
    dev = torch.device("cuda", 0)
    W = H = 256
    add, P0 = 3276, 100_000
    Kmat = synth.intrinsics(W, H).astype(np.float32)
    poses = synth.make_poses(steps + 256, seed=77)
    w2c_all = torch.from_numpy(synth.w2c_from_c2w(poses)).to(dev)
    rsn = O.setup_camera_np(W, H, Kmat, np.eye(4))
    view = torch.from_numpy(rsn["viewmatrix"]).to(dev).unsqueeze(0)
    proj = torch.from_numpy(rsn["projmatrix"]).to(dev).unsqueeze(0)
    campos = torch.from_numpy(rsn["campos"]).to(dev)
    bg, e = torch.zeros(3, device=dev), torch.Tensor([])
    dpix = torch.ones((3, H, W), device=dev) * 1e-3
    raw = synth.make_scene(P0)
    params = {k: torch.from_numpy(np.asarray(v, np.float32)).to(dev) for k, v in raw.items()}
    rng = np.random.default_rng(5)
    nf = min(steps, 64)
    color, depth = _room_frames(poses[:nf], W, H, Kmat, rng, dev)
    masks = torch.zeros((nf, H * W), dtype=torch.bool, device=dev)
    for f in range(nf):
        masks[f, torch.from_numpy(rng.choice(H * W, size=add, replace=False)).to(dev)] = True
    tK = torch.from_numpy(Kmat).to(dev)

    def compute_Hessian(rel_w2c):
        pts = params["means3D"]
        pts4 = torch.cat((pts, torch.ones(pts.shape[0], 1, device=dev)), dim=1)
        tp = (rel_w2c @ pts4.T).T[:, :3].contiguous()
        rot, op, scl = F.normalize(params["unnorm_rotations"]), torch.sigmoid(params["logit_opacities"]), torch.exp(params["log_scales"])
        R, colr, radii, geom, binning, img, dep = ref_C.rasterize_gaussians(
            bg, tp, params["rgb_colors"], op, scl, rot, 1.0, e, view, proj, rsn["tanfovx"], rsn["tanfovy"], H, W, e, 0, campos, False)
        g = ref_C.rasterize_gaussians_backward(bg, tp, radii, params["rgb_colors"], scl, rot, 1.0, e, view, proj, rsn["tanfovx"],
                                               rsn["tanfovy"], dpix, e, 0, campos, geom, R, binning, img, 2)
        _ = int((radii > 0).sum().item())
        return torch.cat([g[3], g[2].reshape(-1, 1)], dim=1)

    cand = torch.from_numpy(synth.make_poses(256, seed=78)).to(dev)
    picks = []
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for s in range(steps):
        f = s % nf
        CX, CY, FX, FY = tK[0][2], tK[1][2], tK[0][0], tK[1][1]
        xg, yg = torch.meshgrid(torch.arange(0, W).cuda().float(), torch.arange(0, H).cuda().float(), indexing="xy")
        xx, yy = ((xg - CX) / FX).reshape(-1), ((yg - CY) / FY).reshape(-1)
        z = depth[f][0].reshape(-1)
        pts4 = torch.cat((torch.stack((xx * z, yy * z, z), dim=-1), torch.ones(H * W, 1).cuda()), dim=1)
        pts = (torch.inverse(w2c_all[f]) @ pts4.T).T[:, :3]
        msd = ((1 * z / ((FX + FY) / 2)) ** 2)[masks[f]]
        cld = torch.cat((pts, torch.permute(color[f], (1, 2, 0)).reshape(-1, 3)), -1)[masks[f]]
        n = cld.shape[0]
        rots = torch.zeros((n, 4), device=dev); rots[:, 0] = 1
        new = {"means3D": cld[:, :3], "rgb_colors": cld[:, 3:6], "unnorm_rotations": rots,
               "logit_opacities": torch.zeros((n, 1), device=dev),
               "log_scales": torch.tile(torch.log(torch.sqrt(msd))[..., None], (1, 3))}
        params = {k: torch.cat((params[k], v), dim=0) for k, v in new.items()}
        vis = len(O.count_visible_points(params["means3D"].cpu().numpy(), poses[s].astype(np.float64), Kmat, W))
        if s % 10 == 9:
            kfs = list(range(0, s + 1, 4))[-32:]
            H_train = sum(compute_Hessian(w2c_all[k]) for k in kfs)
            H_inv = torch.reciprocal(H_train + 0.1)
            sc = [torch.sum(compute_Hessian(torch.linalg.inv(p)) * H_inv).item() for p in cand]
            picks.append(int(np.argmax(sc)))
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    return {"config": f"5 (reference arm, ONE GPU, {steps} steps): torch.cat growth, numpy count_visible_points, reference "
                      "extension driven pose by pose", "steps_per_s": steps / dt, "wall_s": dt, "scoring_rounds": len(picks),
            "picks_head": picks[:8], "visible_last": int(vis), "final_gaussians": int(params["means3D"].shape[0]),
            "host_cores": cores()}


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


def cfg_n1():
    """One get_loss iteration's renders at 200k Gaussians, 256x256: colour render + depth/silhouette
    render, forward and backward -- the fused operator (bins and sorts once) against two calls of this repo's own drop-in
    operator and against two forward + two backward calls of the reference extension."""
    import torch
    import torch.nn.functional as F
    import fisher_nerf_customized_b200 as fb
    from fisher_nerf_customized_b200 import synth
    from oracle import build_ref
    W = H = 256
    P = 200_000
    raw = synth.make_scene(P, seed=21)
    rs = fb.setup_camera(W, H, synth.intrinsics(W, H), np.eye(4, dtype=np.float32))
    w2c = torch.from_numpy(synth.w2c_from_c2w(synth.make_poses(1, seed=4))[0]).cuda()
    p = {k: torch.from_numpy(np.asarray(v, np.float32)).cuda() for k, v in raw.items()}
    pts4 = torch.cat((p["means3D"], torch.ones(P, 1, device="cuda")), dim=1)
    tp = (w2c @ pts4.T).T[:, :3].contiguous()
    rot, op, sc = F.normalize(p["unnorm_rotations"]), torch.sigmoid(p["logit_opacities"]), torch.exp(p["log_scales"])
    ds = fb.get_depth_and_silhouette(tp, torch.eye(4, device="cuda")).contiguous()
    g_im = torch.randn((3, H, W), device="cuda") * 1e-2
    g_ds = torch.randn((3, H, W), device="cuda") * 1e-2
    C = fb._C
    e = torch.Tensor([])

    def fused():
        R, im, ex, radii, geom, inst, img, d = C.rasterize_gaussians2(rs.bg, tp, p["rgb_colors"], ds, op, sc, rot, 1.0, rs.viewmatrix,
                                                                     rs.projmatrix, rs.tanfovx, rs.tanfovy, H, W, rs.campos, False)
        return C.rasterize_gaussians_backward2(rs.bg, tp, radii, p["rgb_colors"], ds, sc, rot, 1.0, rs.viewmatrix, rs.projmatrix,
                                               rs.tanfovx, rs.tanfovy, g_im, g_ds, rs.campos, geom, R, inst, img)

    def two(CC):
        out = []
        for col, g in ((p["rgb_colors"], g_im), (ds, g_ds)):
            R, im, radii, geom, inst, img, d = CC.rasterize_gaussians(rs.bg, tp, col, op, sc, rot, 1.0, e, rs.viewmatrix, rs.projmatrix,
                                                                     rs.tanfovx, rs.tanfovy, H, W, e, 0, rs.campos, False)
            out.append(CC.rasterize_gaussians_backward(rs.bg, tp, radii, col, sc, rot, 1.0, e, rs.viewmatrix, rs.projmatrix, rs.tanfovx,
                                                       rs.tanfovy, g, e, 0, rs.campos, geom, R, inst, img, 1))
        return out
    out = {"config": "N1: the two renders of one get_loss iteration (colour + depth/silhouette), forward + backward, 200k Gaussians, 256x256",
           "fused_ms": ev_time(fused, 30), "two_dropin_calls_ms": ev_time(lambda: two(C), 30)}
    ref = build_ref.import_ref("ref_dgr")
    if ref is not None:
        out["reference_extension_two_calls_ms"] = ev_time(lambda: two(ref), 10)
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", default="1,2,4,5,5a")
    ap.add_argument("--loop-steps", type=int, default=200)
    ap.add_argument("--ref-steps", type=int, default=30)
    args = ap.parse_args()
    from fisher_nerf_customized_b200 import _lib
    _lib.lib()
    todo = {"1": cfg1, "2": cfg2, "3": cfg3_variants, "4": cfg4, "5": lambda: cfg5(args.loop_steps), "5a": cfg5a,
            "5ref": lambda: cfg5_reference(args.ref_steps), "n1": cfg_n1}
    for k in args.only.split(","):
        r = todo[k.strip()]()
        if r is not None:
            print(json.dumps(r), flush=True)


if __name__ == "__main__":
    main()
