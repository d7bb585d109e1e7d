"""Generate the golden fixtures under tests/golden/ from THE REFERENCE ITSELF.

  python tests/golden/make_golden.py gpu   # on the GPU box: runs the unmodified reference CUDA extensions
                                           # (oracle/_ref/ref_dgr, ref_knn; built by oracle/build_ref.py from
                                           # /root/reference) on seeded inputs -> gpurun_out/golden/*.npz
  python tests/golden/make_golden.py blocks  # in the build container: executes the reference's own t_opt_blocks /
                                           # d_opt_blocks / topt_score_from_diags / dopt_score_from_diags /
                                           # compute_H_train_blocks method bodies (models/SLAM/gaussian_object.py,
                                           # taken by AST: the module imports the whole SLAM stack) on seeded blocks
  python tests/golden/make_golden.py pointcloud  # in the build container: the reference's get_pointcloud /
                                           # initialize_new_params bodies (models/SLAM/gaussian.py, by AST) on CPU
                                           # torch (`.cuda()` made a no-op) on seeded RGB-D frames
  python tests/golden/make_golden.py maskcount  # in the build container: the reference's count_gaussians_vs_mask body
                                           # (models/SLAM/gaussian_object.py:359-428, by AST) on CPU torch
  python tests/golden/make_golden.py cpu   # in the build container: executes the reference's own numpy
                                           # count_visible_points, taken from /root/reference/tester_navigator.py
                                           # by AST (the module itself imports habitat and cannot be imported)

Inputs are NOT stored: they are regenerated from the seeds by tests/common.py / synth.py; each fixture
records the seed and shape it was made with.  The fixtures pin the CPU oracle (tests/test_oracle.py) and
are compared with the CUDA path in the -m gpu tests.
"""
import ast
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

RASTER_CASES = [   # name, P, W, H, seed, identity view?
    ("raster_a", 2500, 64, 64, 101, True),
    ("raster_b", 4000, 96, 80, 102, False),
]
KNN_CASES = [("knn_a", 5000, 201), ("knn_b", 3, 202), ("knn_c", 4, 203)]
VIS_CASES = [("vis_a", 200_000, 16, 301)]


def _img_state(img, HW):
    """ImageState::fromChunk (DGR/cuda_rasterizer/rasterizer_impl.cu:177-185): accum_alpha f32[N],
    n_contrib u32[N], ranges uint2[N], each 128-byte aligned from the (aligned) chunk base."""
    import torch
    base = img.data_ptr()

    def al(p):
        return (p + 127) // 128 * 128
    o0 = al(base) - base
    accum = img[o0:o0 + 4 * HW].view(torch.float32)
    o1 = al(base + o0 + 4 * HW) - base
    ncon = img[o1:o1 + 4 * HW].view(torch.int32)
    o2 = al(base + o1 + 4 * HW) - base
    ranges = img[o2:o2 + 8 * HW].view(torch.int32)
    return accum.cpu().numpy(), ncon.cpu().numpy().view(np.uint32), ranges.cpu().numpy().view(np.uint32)


def _walk(buf):
    """obtain() of rasterizer_impl.h:21-28: consecutive arrays, each aligned to 128 bytes of ABSOLUTE address."""
    base = buf.data_ptr()
    state = {"p": base}

    def take(nbytes):
        o = (state["p"] + 127) // 128 * 128 - base
        state["p"] = base + o + nbytes
        return buf[o:o + nbytes]
    return take


def _geom_state(geom, P):
    """GeometryState::fromChunk (rasterizer_impl.cu:155-170): depths f32[P], clamped bool[3P], internal_radii i32[P],
    means2D float2[P], cov3D f32[6P], conic_opacity float4[P], rgb f32[3P], tiles_touched u32[P], ..."""
    import torch
    take = _walk(geom)
    depths = take(4 * P).view(torch.float32)
    take(3 * P)
    take(4 * P)
    means2D = take(8 * P).view(torch.float32).reshape(P, 2)
    cov3D = take(24 * P).view(torch.float32).reshape(P, 6)
    conic_opacity = take(16 * P).view(torch.float32).reshape(P, 4)
    take(12 * P)
    tiles_touched = take(4 * P).view(torch.int32)
    return (depths.cpu().numpy(), means2D.cpu().numpy(), cov3D.cpu().numpy(), conic_opacity.cpu().numpy(),
            tiles_touched.cpu().numpy().view(np.uint32))


def _binning_state(binning, R):
    """BinningState::fromChunk (rasterizer_impl.cu:181-194): point_list u32[R], point_list_unsorted u32[R],
    point_list_keys u64[R], point_list_keys_unsorted u64[R], ..."""
    import torch
    take = _walk(binning)
    point_list = take(4 * R).view(torch.int32)
    take(4 * R)
    keys = take(8 * R).view(torch.int64)
    return point_list.cpu().numpy().view(np.uint32), keys.cpu().numpy().view(np.uint64)


def gpu(out_dir):
    import torch
    from oracle import build_ref, oracle as O
    from tests import common
    from fisher_nerf_customized_b200 import synth
    os.makedirs(out_dir, exist_ok=True)
    dgr, knn = build_ref.import_ref("ref_dgr"), build_ref.import_ref("ref_knn")
    assert dgr is not None and knn is not None, "oracle/_ref not built"
    e = torch.Tensor([])
    for name, P, W, H, seed, ident in RASTER_CASES:
        sc = common.random_view_scene(P, seed, W, H)
        w2c = np.eye(4, dtype=np.float32) if ident else common.look_at_w2c((0.3, -0.2, -0.5), (0.1, 0.0, 3.0))
        rs = O.setup_camera_np(W, H, common.K_for(W, H), w2c)
        t = {k: torch.from_numpy(v).cuda() for k, v in sc.items()}
        bg = torch.zeros(3, device="cuda")
        view = torch.from_numpy(rs["viewmatrix"]).cuda().unsqueeze(0)
        proj = torch.from_numpy(rs["projmatrix"]).cuda().unsqueeze(0)
        campos = torch.from_numpy(rs["campos"]).cuda()
        R, color, radii, geom, binning, img, depth = dgr.rasterize_gaussians(
            bg, t["means3D"], t["rgb_colors"], t["opacities"], t["scales"], t["rotations"], 1.0, e, view, proj,
            rs["tanfovx"], rs["tanfovy"], H, W, e, 0, campos, False)
        final_T, n_contrib, ranges = _img_state(img, H * W)
        tiles = ((W + 15) // 16) * ((H + 15) // 16)
        out = dict(P=P, W=W, H=H, seed=seed, ident=ident, num_rendered=R, color=color.cpu().numpy(),
                   radii=radii.cpu().numpy(), depth=depth.cpu().numpy(), final_T=final_T, n_contrib=n_contrib,
                   ranges=ranges[:2 * tiles].reshape(tiles, 2))
        # the reference's own geometry and binning state (only rows of Gaussians with radii > 0 are ever written)
        vis = (radii > 0).cpu().numpy()
        depths, means2D, cov3D, conic_opacity, tiles_touched = _geom_state(geom, P)
        point_list, point_keys = _binning_state(binning, R)
        out.update(point_list=point_list, point_list_keys=point_keys, tiles_touched=np.where(vis, tiles_touched, 0),
                   depths=np.where(vis, depths, 0), means2D=np.where(vis[:, None], means2D, 0),
                   conic_opacity=np.where(vis[:, None], conic_opacity, 0))
        rng = np.random.default_rng(seed)
        dpix1 = (rng.normal(size=(3, H, W)) * 1e-3).astype(np.float32)
        for power, dpix in ((1, dpix1), (2, np.full((3, H, W), 1e-3, np.float32))):
            g = dgr.rasterize_gaussians_backward(bg, t["means3D"], radii, t["rgb_colors"], t["scales"], t["rotations"],
                                                 1.0, e, view, proj, rs["tanfovx"], rs["tanfovy"],
                                                 torch.from_numpy(dpix).cuda(), e, 0, campos, geom, R, binning, img, power)
            names = ("dL_dmeans2D", "dL_dcolors", "dL_dopacity", "dL_dmeans3D", "dL_dcov3D", "dL_dsh", "dL_dscales",
                     "dL_drotations")
            for n, v in zip(names, g):
                if n != "dL_dsh":
                    out[f"p{power}_{n}"] = v.cpu().numpy()
        np.savez_compressed(os.path.join(out_dir, name + ".npz"), **out)
        print(name, "R", R, "visible", int((radii > 0).sum()))
    for name, P, seed in KNN_CASES:
        pts = synth.make_points(P, seed=seed)
        d = knn.distCUDA2(torch.from_numpy(pts).cuda()).cpu().numpy()
        np.savez_compressed(os.path.join(out_dir, name + ".npz"), P=P, seed=seed, dist2=d)
        print(name, d[:3])


def _reference_count_visible_points():
    src = open("/root/reference/tester_navigator.py").read()
    tree = ast.parse(src)
    fn = next(n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name == "count_visible_points")
    ns = {"np": np}
    exec(compile(ast.Module(body=[fn], type_ignores=[]), "tester_navigator.py", "exec"), ns)
    return ns["count_visible_points"]


def cpu(out_dir):
    from fisher_nerf_customized_b200 import synth
    os.makedirs(out_dir, exist_ok=True)
    ref_fn = _reference_count_visible_points()
    for name, N, K, seed in VIS_CASES:
        pts = synth.make_points(N, seed=seed)
        c2w = synth.make_poses(K, seed=seed + 1).astype(np.float64)
        Kmat = synth.intrinsics(256, 256)
        counts = np.array([len(ref_fn(pts, c2w[k], Kmat, 256)) for k in range(K)], dtype=np.int64)
        first = ref_fn(pts, c2w[0], Kmat, 256)
        np.savez_compressed(os.path.join(out_dir, name + ".npz"), N=N, K=K, seed=seed, counts=counts,
                            first_pose_checksum=np.float64(first.astype(np.float64).sum()), first_pose_n=len(first))
        print(name, counts)


BLOCK_CASES = [("blocks_a", 48, 11, 401), ("blocks_b", 33, 7, 402), ("blocks_c", 20, 3, 403)]


def block_inputs(n, d, seed):
    """Seeded rank-deficient PSD blocks like a few probes x a few keyframes produce (shared with the tests)."""
    rng = np.random.default_rng(seed)
    v = rng.standard_normal((n, 4, d)).astype(np.float32)
    w = rng.standard_normal((n, 2, d)).astype(np.float32)
    Hm = np.einsum("nri,nrj->nij", v, v).astype(np.float32)
    J = np.einsum("nri,nrj->nij", w, w).astype(np.float32)
    return Hm, J


def _reference_methods(names):
    import torch
    src = open("/root/reference/models/SLAM/gaussian_object.py").read()
    ns = {"torch": torch, "np": np}
    for node in ast.walk(ast.parse(src)):
        if isinstance(node, ast.FunctionDef) and node.name in names and node.name not in ns:
            exec(compile(ast.Module(body=[node], type_ignores=[]), "gaussian_object.py", "exec"), ns)
    return [ns[n] for n in names]


def blocks(out_dir):
    import torch
    t_opt, d_opt, t_diag, d_diag, h_train = _reference_methods(
        ["t_opt_blocks", "d_opt_blocks", "topt_score_from_diags", "dopt_score_from_diags", "compute_H_train_blocks"])
    for name, n, d, seed in BLOCK_CASES:
        Hm, J = block_inputs(n, d, seed)
        out = dict(n=n, d=d, seed=seed)
        for lam in (1e-2, 1e-4):
            # the reference's own arithmetic: FP32 (as the callers run it) and the same method on FP64 tensors
            for tag, dt in (("f32", torch.float32), ("f64", torch.float64)):
                a, b = torch.from_numpy(Hm).to(dt), torch.from_numpy(J).to(dt)
                out[f"topt_{tag}_{lam:g}"] = float(t_opt(None, a, b, lam))
                out[f"dopt_{tag}_{lam:g}"] = float(d_opt(None, a, b, lam))
        hd, jd = np.einsum("bii->bi", Hm).reshape(-1), np.einsum("bii->bi", J).reshape(-1)
        out["topt_diag"] = float(t_diag(None, torch.from_numpy(hd), torch.from_numpy(jd), 1e-6))
        out["dopt_diag"] = float(d_diag(None, torch.from_numpy(hd), torch.from_numpy(jd), 1e-6))

        # compute_H_train_blocks with three keyframes of different visible counts (its positional truncation)
        class Stub:
            keyframe_list = [dict(est_w2c=k) for k in range(3)]

            def estimate_block_JtJ(self, w2c, K=2, **kw):
                m = [n, n - 5, n - 2][w2c]
                return torch.from_numpy(Hm[:m] * np.float32(w2c + 1)), torch.arange(m) * (w2c + 1)
        Hsum, vis = h_train(Stub(), K=2)
        out["htrain_blocks"], out["htrain_vis"] = Hsum.numpy(), vis.numpy()
        np.savez_compressed(os.path.join(out_dir, name + ".npz"), **out)
        print(name, {k: v for k, v in out.items() if isinstance(v, float)})


PCD_CASES = [("pcd_a", 48, 64, 1, 501), ("pcd_b", 48, 64, 2, 502), ("pcd_c", 40, 56, 4, 503)]


def frame_inputs(H, W, seed):
    """Seeded RGB-D frame, intrinsics, w2c and a non-presence-like mask (shared with the tests)."""
    rng = np.random.default_rng(seed)
    color = rng.random((3, H, W)).astype(np.float32)
    depth = (0.5 + 4.0 * rng.random((1, H, W))).astype(np.float32)
    depth[0, rng.random((H, W)) < 0.05] = 0.0                          # holes, as in sensor depth
    K = np.array([[W / 2.0, 0, W / 2.0 - 0.5], [0, W / 2.0 + 3.0, H / 2.0 + 0.25], [0, 0, 1]], np.float32)
    a = 0.3 + 0.1 * seed % 7
    w2c = np.eye(4, dtype=np.float32)
    w2c[:3, :3] = np.array([[np.cos(a), 0, np.sin(a)], [0, 1, 0], [-np.sin(a), 0, np.cos(a)]], np.float32)
    w2c[:3, 3] = rng.normal(0, 1, 3).astype(np.float32)
    mask = (rng.random(H * W) < 0.3) & (depth.reshape(-1) > 0.01)
    return color, depth, K, w2c, mask


def pointcloud(out_dir):
    import torch
    import torch.nn.functional as F
    from einops import rearrange
    src = open("/root/reference/models/SLAM/gaussian.py").read()
    class TorchOnCpu:                                                  # `device="cuda"` factory calls -> CPU
        def __getattr__(self, k):
            v = getattr(torch, k)
            if k in ("zeros", "ones", "eye", "tensor", "arange", "empty"):
                return lambda *a, **kw: v(*a, **{x: y for x, y in kw.items() if x != "device"})
            return v
    ns = {"torch": TorchOnCpu(), "np": np, "F": F, "rearrange": rearrange}
    for node in ast.parse(src).body:
        if isinstance(node, ast.FunctionDef) and node.name in ("get_pointcloud", "initialize_new_params"):
            exec(compile(ast.Module(body=[node], type_ignores=[]), "gaussian.py", "exec"), ns)
    torch.Tensor.cuda = lambda self, *a, **k: self                    # the reference's body calls .cuda() on everything
    for name, H, W, ds, seed in PCD_CASES:
        color, depth, K, w2c, mask = frame_inputs(H, W, seed)
        t = lambda x: torch.from_numpy(x)
        cld, msd = ns["get_pointcloud"](t(color), t(depth), t(K), t(w2c), mask=t(mask), compute_mean_sq_dist=True,
                                        downsample=ds)
        cld_all = ns["get_pointcloud"](t(color), t(depth), t(K), t(w2c), downsample=ds)
        cld_cam = ns["get_pointcloud"](t(color), t(depth), t(K), t(w2c), transform_pts=False, downsample=ds)
        pos = msd > 0
        newp = ns["initialize_new_params"](cld[pos], msd[pos], isotropic=(ds == 2))
        np.savez_compressed(os.path.join(out_dir, name + ".npz"), H=H, W=W, ds=ds, seed=seed, cld=cld.numpy(),
                            msd=msd.numpy(), cld_all=cld_all.numpy(), cld_cam=cld_cam.numpy(),
                            **{"new_" + k: v.detach().numpy() for k, v in newp.items()})
        print(name, cld.shape, cld_all.shape, float(msd.min()), float(msd.max()))


def mask_inputs(N, H, W, seed):
    """Seeded Gaussians (means, opacity logits), a pose, intrinsics and an object mask (shared with the tests)."""
    from tests import common
    rng = np.random.default_rng(seed)
    means = np.c_[rng.uniform(-3, 3, N), rng.uniform(-2, 2, N), rng.uniform(-1.0, 6.0, N)].astype(np.float32)
    logit = rng.normal(-2.0, 3.0, size=(N, 1)).astype(np.float32)
    w2c = common.look_at_w2c((0.3, -0.1, -0.5), (0.1, 0.0, 2.5))
    K = np.array([[0.8 * W, 0, W / 2.0], [0, 0.8 * W, H / 2.0], [0, 0, 1]], np.float32)
    yy, xx = np.mgrid[0:H, 0:W]
    mask = ((xx - 0.55 * W) ** 2 + (yy - 0.45 * H) ** 2 < (0.3 * H) ** 2) | (rng.random((H, W)) < 0.05)
    return means, logit, w2c, K, mask


MASK_CASES = [("mask_a", 20000, 48, 64, 401), ("mask_b", 7, 16, 16, 402)]


def maskcount(out_dir):
    import torch
    fn, = _reference_methods(["count_gaussians_vs_mask"])
    for name, N, H, W, seed in MASK_CASES:
        means, logit, w2c, K, mask = mask_inputs(N, H, W, seed)
        params = {"means3D": torch.from_numpy(means), "logit_opacities": torch.from_numpy(logit)}
        r = fn(params, torch.from_numpy(w2c), torch.from_numpy(K), torch.from_numpy(mask), 0.01)
        np.savez_compressed(os.path.join(out_dir, name + ".npz"), n_in=r["in"], n_out=r["out"], active=r["active"],
                            idx_in=r["idx_in"].numpy(), idx_out=r["idx_out"].numpy(), N=N, H=H, W=W, seed=seed)
        print(name, r["in"], r["out"], r["active"])


if __name__ == "__main__":
    mode = sys.argv[1]
    if mode == "maskcount":
        maskcount(HERE)
    elif mode == "pointcloud":
        pointcloud(HERE)
    elif mode == "blocks":
        blocks(HERE)
    elif mode == "gpu":
        gpu(os.path.join(ROOT, "gpurun_out", "golden"))
    else:
        cpu(HERE)
