"""Python face of the CPU oracle (oracle/fisher_oracle.cpp) -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / reference legs may import this
module, and only as the checker.  The product package never does.

Parity status: the reference has no tests or fixtures for this path; the oracle
is pinned against outputs of the reference CUDA extension itself (rebuilt by oracle/build_ref.py,
run by tests/golden/make_golden.py on the GPU box, fixtures committed under tests/golden/).

Everything is numpy in / numpy out.  Function names follow the reference:
  rasterize_forward / rasterize_backward   DGR/rasterize_points.cu:35-196
  compute_Hessian / pose_eval               models/SLAM/gaussian.py:1503-1570, 1354-1375
  distCUDA2                                 KNN/simple_knn.cu:186-222
  count_visible_points                      tester_navigator.py:147-180
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "fisher_oracle.cpp")
LIB = os.path.join(HERE, "liboracle.so")


def build(force=False):
    """g++ -O2 -ffp-contract=off (the file spells out the reference's FMA placement itself)."""
    if force or not os.path.exists(LIB) or os.path.getmtime(LIB) < os.path.getmtime(SRC):
        subprocess.run(["g++", "-O2", "-ffp-contract=off", "-std=c++17", "-shared", "-fPIC", "-o", LIB, SRC],
                       check=True)
    return LIB


class Camera(C.Structure):   # field order of fb_camera / oracle Camera
    _fields_ = [("W", C.c_int), ("H", C.c_int), ("tanfovx", C.c_float), ("tanfovy", C.c_float),
                ("scale_modifier", C.c_float), ("D", C.c_int), ("M", C.c_int), ("prefiltered", C.c_int),
                ("bg", C.c_float * 3), ("campos", C.c_float * 3), ("view", C.c_float * 16), ("proj", C.c_float * 16)]


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(LIB)
        vp = C.c_void_p
        L.orc_forward.restype = vp
        L.orc_forward.argtypes = [vp, C.c_int] + [vp] * 7
        L.orc_num_rendered.restype = C.c_longlong
        L.orc_num_rendered.argtypes = [vp]
        L.orc_prefiltered_trap.restype = C.c_int
        L.orc_prefiltered_trap.argtypes = [vp]
        L.orc_export.restype = None
        L.orc_export.argtypes = [vp] * 16
        L.orc_backward.restype = C.c_int
        L.orc_backward.argtypes = [vp, vp, C.c_int] + [vp] * 12
        L.orc_free.restype = None
        L.orc_free.argtypes = [vp]
        L.orc_mark_visible.restype = None
        L.orc_mark_visible.argtypes = [C.c_int, vp, vp, vp]
        L.orc_knn3.restype = None
        L.orc_knn3.argtypes = [C.c_int, vp, vp]
        L.orc_count_visible.restype = None
        L.orc_count_visible.argtypes = [C.c_int, vp, C.c_int, vp, C.c_double, C.c_double, C.c_double, C.c_double,
                                        C.c_int, C.c_int, vp]
        _lib = L
    return _lib


def _f32(a):
    return None if a is None else np.ascontiguousarray(np.asarray(a, dtype=np.float32))


def _p(a):
    return None if a is None or a.size == 0 else a.ctypes.data_as(C.c_void_p)


def make_camera(W, H, tanfovx, tanfovy, viewmatrix, projmatrix, bg=(0, 0, 0), scale_modifier=1.0, sh_degree=0, M=0,
                campos=(0, 0, 0), prefiltered=False):
    """viewmatrix / projmatrix: the 16 floats exactly as the reference passes them (transposed matrices)."""
    cam = Camera()
    cam.W, cam.H, cam.tanfovx, cam.tanfovy = int(W), int(H), float(tanfovx), float(tanfovy)
    cam.scale_modifier, cam.D, cam.M, cam.prefiltered = float(scale_modifier), int(sh_degree), int(M), int(prefiltered)
    cam.bg[:] = [float(x) for x in bg]
    cam.campos[:] = [float(x) for x in campos]
    cam.view[:] = [float(x) for x in np.asarray(viewmatrix, dtype=np.float32).reshape(-1)]
    cam.proj[:] = [float(x) for x in np.asarray(projmatrix, dtype=np.float32).reshape(-1)]
    return cam


def setup_camera_np(w, h, k, w2c, near=0.01, far=100):
    """numpy restatement of setup_camera (models/SLAM/utils/recon_helpers.py:4-32): returns the dict of
    GaussianRasterizationSettings fields with viewmatrix/projmatrix as the transposed float32 matrices."""
    fx, fy, cx, cy = float(k[0][0]), float(k[1][1]), float(k[0][2]), float(k[1][2])
    w2c = np.asarray(w2c, dtype=np.float32)
    cam_center = np.linalg.inv(w2c)[:3, 3]
    opengl_proj = np.array([[2 * fx / w, 0.0, -(w - 2 * cx) / w, 0.0],
                            [0.0, 2 * fy / h, -(h - 2 * cy) / h, 0.0],
                            [0.0, 0.0, far / (far - near), -(far * near) / (far - near)],
                            [0.0, 0.0, 1.0, 0.0]], dtype=np.float32)
    view_t = np.ascontiguousarray(w2c.T)
    full_proj = (view_t @ opengl_proj.T).astype(np.float32)
    return dict(image_height=h, image_width=w, tanfovx=w / (2 * fx), tanfovy=h / (2 * fy),
                bg=np.zeros(3, np.float32), scale_modifier=1.0, viewmatrix=view_t, projmatrix=full_proj, sh_degree=0,
                campos=cam_center.astype(np.float32), prefiltered=False)


def camera_from_settings(rs, M=0):
    return make_camera(rs["image_width"], rs["image_height"], rs["tanfovx"], rs["tanfovy"], rs["viewmatrix"],
                       rs["projmatrix"], rs["bg"], rs["scale_modifier"], rs["sh_degree"], M, rs["campos"],
                       rs["prefiltered"])


class Forward:
    """Result of one oracle forward; holds the native state for backward()."""

    def __init__(self, cam, means3D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                 cov3D_precomp=None):
        L = lib()
        self.cam = cam
        self.P = P = int(np.asarray(means3D).shape[0])
        self._in = [_f32(means3D), _f32(np.asarray(opacities).reshape(-1)), _f32(shs), _f32(colors_precomp),
                    _f32(scales), _f32(rotations), _f32(cov3D_precomp)]
        self._h = L.orc_forward(C.byref(cam), P, *[_p(a) for a in self._in])
        if not self._h:
            raise ValueError("oracle: invalid argument combination")
        self.num_rendered = int(L.orc_num_rendered(self._h))
        self.prefiltered_trap = bool(L.orc_prefiltered_trap(self._h))
        W, H = cam.W, cam.H
        tiles = ((W + 15) // 16) * ((H + 15) // 16)
        R = self.num_rendered
        self.radii = np.zeros(P, np.int32)
        self.tiles_touched = np.zeros(P, np.uint32)
        self.depths = np.zeros(P, np.float32)
        self.means2D = np.zeros((P, 2), np.float32)
        self.cov3D = np.zeros((P, 6), np.float32)
        self.conic_opacity = np.zeros((P, 4), np.float32)
        self.rgb = np.zeros((P, 3), np.float32)
        self.clamped = np.zeros((P, 3), np.uint8)
        self.keys = np.zeros(R, np.uint64)
        self.point_list = np.zeros(R, np.uint32)
        self.ranges = np.zeros((tiles, 2), np.uint32)
        self.final_T = np.zeros(H * W, np.float32)
        self.n_contrib = np.zeros(H * W, np.uint32)
        self.color = np.zeros((3, H, W), np.float32)
        self.depth = np.zeros((1, H, W), np.float32)
        L.orc_export(self._h, *[_p(a) for a in (self.radii, self.tiles_touched, self.depths, self.means2D, self.cov3D,
                                                self.conic_opacity, self.rgb, self.clamped, self.keys, self.point_list,
                                                self.ranges, self.final_T, self.n_contrib, self.color, self.depth)])

    def backward(self, dL_dpix, power=1, f64=False):
        """Returns dict of the reference's gradient tensors (float32; with f64=True also the FP64 sums
        of means3D / opacity / scales / rotations under the keys '*_f64')."""
        L = lib()
        P = self.P
        g = dict(dL_dmeans2D=np.zeros((P, 3), np.float32), dL_dconic=np.zeros((P, 2, 2), np.float32),
                 dL_dopacity=np.zeros((P, 1), np.float32), dL_dcolors=np.zeros((P, 3), np.float32),
                 dL_dmeans3D=np.zeros((P, 3), np.float32), dL_dcov3D=np.zeros((P, 6), np.float32),
                 dL_dscales=np.zeros((P, 3), np.float32), dL_drotations=np.zeros((P, 4), np.float32))
        d = dict(dL_dmeans3D_f64=np.zeros((P, 3)), dL_dopacity_f64=np.zeros((P, 1)), dL_dscales_f64=np.zeros((P, 3)),
                 dL_drotations_f64=np.zeros((P, 4)))
        dpix = _f32(dL_dpix)
        assert dpix.shape == (3, self.cam.H, self.cam.W)
        rc = L.orc_backward(self._h, _p(dpix), int(power), *[_p(g[k]) for k in
                            ("dL_dmeans2D", "dL_dconic", "dL_dopacity", "dL_dcolors", "dL_dmeans3D", "dL_dcov3D",
                             "dL_dscales", "dL_drotations")],
                            *[_p(d[k]) if f64 else None for k in
                              ("dL_dmeans3D_f64", "dL_dopacity_f64", "dL_dscales_f64", "dL_drotations_f64")])
        if rc != 0:
            raise NotImplementedError("oracle backward: SH degree > 0 is not restated")
        if f64:
            g.update(d)
        return g

    def close(self):
        if self._h:
            lib().orc_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def mark_visible(means3D, viewmatrix):
    m = _f32(means3D)
    v = _f32(np.asarray(viewmatrix).reshape(-1))
    out = np.zeros(m.shape[0], np.uint8)
    lib().orc_mark_visible(m.shape[0], _p(m), _p(v), _p(out))
    return out.astype(bool)


def distCUDA2(points):
    p = _f32(points)
    out = np.zeros(p.shape[0], np.float32)
    lib().orc_knn3(p.shape[0], _p(p), _p(out))
    return out


def count_visible_points(global_pts, c2w, intrinsics, img_size):
    """numpy restatement of count_visible_points (tester_navigator.py:147-180), same operations in
    the same order so that numpy's float64 promotion and BLAS matmul are the ones the reference uses:
    inv(c2w); (R @ pts.T + t).T; keep z > 0; u = fx*x/z + cx, v = fy*y/z + cy; keep 0<=u<W, 0<=v<H.
    Returns the visible subset of global_pts, input order."""
    fx, fy, cx, cy = intrinsics[0, 0], intrinsics[1, 1], intrinsics[0, 2], intrinsics[1, 2]
    H = W = img_size
    w2c = np.linalg.inv(c2w)
    cam_pts = (w2c[:3, :3] @ global_pts.T + w2c[:3, 3:4]).T
    front = cam_pts[:, 2] > 0
    cam_pts, kept = cam_pts[front], global_pts[front]
    u = fx * cam_pts[:, 0] / cam_pts[:, 2] + cx
    v = fy * cam_pts[:, 1] / cam_pts[:, 2] + cy
    return kept[(u >= 0) & (u < W) & (v >= 0) & (v < H)]


def count_visible_counts_c(global_pts, c2w_list, intrinsics, img_size):
    """C port of the same test for many poses (counts only); used where numpy would be slow."""
    p = _f32(global_pts)
    w2c = np.ascontiguousarray(np.stack([np.linalg.inv(np.asarray(m)) for m in c2w_list]).astype(np.float64))
    out = np.zeros(w2c.shape[0], np.int32)
    k = np.asarray(intrinsics)
    lib().orc_count_visible(p.shape[0], _p(p), w2c.shape[0], w2c.ctypes.data_as(C.c_void_p), float(k[0, 0]),
                            float(k[1, 1]), float(k[0, 2]), float(k[1, 2]), int(img_size), int(img_size), _p(out))
    return out


# ---- caller arithmetic (models/SLAM/gaussian.py) ----------------------------------------------------

def transform_points_f32(rel_w2c, pts):
    """(rel_w2c @ [pts|1].T).T[:, :3] in float32 (gaussian.py:1523-1527).  torch evaluates this with
    an FP32 GEMM over k = 4; a k-sequential FMA accumulation is fma(m3,1,fma(m2,z,fma(m1,y,m0*x)))."""
    m = np.asarray(rel_w2c, dtype=np.float32)
    p = np.asarray(pts, dtype=np.float32)
    out = np.empty_like(p)
    for r in range(3):
        acc = (m[r, 0] * p[:, 0]).astype(np.float32)
        # fused multiply-adds emulated in float64 (exact product, one rounding)
        acc = (m[r, 1].astype(np.float64) * p[:, 1].astype(np.float64) + acc.astype(np.float64)).astype(np.float32)
        acc = (m[r, 2].astype(np.float64) * p[:, 2].astype(np.float64) + acc.astype(np.float64)).astype(np.float32)
        out[:, r] = acc + m[r, 3]
    return out


def compute_Hessian(rs, params, rel_w2c, C_H=4, grad_scale=1e-3, f64=True):
    """cur_H[P, C_H] of one pose: compute_Hessian(return_points=True), gaussian.py:1503-1570 (C_H = 4:
    means3D grad (3) + opacity grad (1)); gaussian_object.py:2024-2032 for C_H = 11 (+ scales (3) +
    rotations (4)).  rs: settings dict from setup_camera_np (identity view matrix in the reference);
    params: dict(means3D, rgb_colors, rotations (normalised), opacities, scales) in activated form."""
    cam = camera_from_settings(rs)
    pts_cam = transform_points_f32(rel_w2c, params["means3D"])
    fwd = Forward(cam, pts_cam, params["opacities"], colors_precomp=params["rgb_colors"], scales=params["scales"],
                  rotations=params["rotations"])
    dpix = np.full((3, cam.H, cam.W), grad_scale, np.float32)
    g = fwd.backward(dpix, power=2, f64=f64)
    sfx = "_f64" if f64 else ""
    cols = [g["dL_dmeans3D" + sfx], g["dL_dopacity" + sfx]]
    if C_H == 11:
        cols += [g["dL_dscales" + sfx], g["dL_drotations" + sfx]]
    H = np.concatenate([c.reshape(fwd.P, -1) for c in cols], axis=1)
    vis = fwd.radii > 0
    fwd.close()
    return H, vis


def compute_H_train(rs, params, keyframe_w2c, C_H=4, grad_scale=1e-3):
    """gaussian.py:1338-1348: sum of per-keyframe Hessians."""
    H = None
    for w2c in keyframe_w2c:
        h, _ = compute_Hessian(rs, params, w2c, C_H, grad_scale)
        H = h if H is None else H + h
    return H


def pose_eval(rs, params, poses_w2c, H_train_inv, C_H=4, grad_scale=1e-3):
    """gaussian.py:1362-1370: scores[k] = sum(cur_H_k * H_train_inv)."""
    out = []
    for w2c in poses_w2c:
        h, _ = compute_Hessian(rs, params, w2c, C_H, grad_scale)
        out.append(float(np.sum(h * np.asarray(H_train_inv, dtype=np.float64))))
    return np.array(out)


# ---- POp-GS block form (models/SLAM/gaussian_object.py), numpy restatement ---------------------------------------

def block_outer(G_rows, K=None):
    """estimate_block_JtJ's accumulation (gaussian_object.py:2160-2174) for a list of per-probe gradient matrices
    G_rows[k] = Gv [Nv,d] (visible rows, columns means|opacity|rotations|scales): sum_k Gv (x) Gv, divided by K.
    FP32 with the reference's rounding points (product rounded, then added; one division at the end)."""
    H = None
    for Gv in G_rows:
        Gv = np.asarray(Gv, dtype=np.float32)
        o = (Gv[:, :, None] * Gv[:, None, :]).astype(np.float32)
        H = o if H is None else (H + o).astype(np.float32)
    return (H / np.float32(len(G_rows) if K is None else K)).astype(np.float32)


def t_opt_blocks(Hm_blocks, J_blocks, lam=1e-6):
    """t_opt_blocks (gaussian_object.py:1720-1724): -sum_b trace((Hm_b + J_b + lam I)^-1).  The matrix is formed in
    FP32 as the reference forms it and inverted in FP64 (LAPACK LU with partial pivoting, as torch.linalg.inv)."""
    Hm, J = np.asarray(Hm_blocks, np.float32), np.asarray(J_blocks, np.float32)
    if Hm.shape[0] == 0:
        return 0.0
    I = np.eye(Hm.shape[-1], dtype=np.float32)
    M = ((Hm + J).astype(np.float32) + np.float32(lam) * I).astype(np.float32)
    return float(-np.einsum("bii->", np.linalg.inv(M.astype(np.float64))))


def d_opt_blocks(Hm_blocks, J_blocks, lam=1e-6):
    """d_opt_blocks (gaussian_object.py:1726-1732): sum_b logabsdet(Hm_b + lam I + J_b) - logabsdet(Hm_b + lam I)."""
    Hm, J = np.asarray(Hm_blocks, np.float32), np.asarray(J_blocks, np.float32)
    if Hm.shape[0] == 0:
        return 0.0
    I = np.eye(Hm.shape[-1], dtype=np.float32)
    M0 = (Hm + np.float32(lam) * I).astype(np.float32)
    M1 = (M0 + J).astype(np.float32)
    return float((np.linalg.slogdet(M1.astype(np.float64))[1] - np.linalg.slogdet(M0.astype(np.float64))[1]).sum())


def compute_H_train_blocks(per_keyframe):
    """compute_H_train_blocks (gaussian_object.py:1572-1586) over [(H_blocks, vis_idx), ...]: positional truncation
    to the smaller visible count."""
    Hm, vis_ref = None, None
    for Hb, vis in per_keyframe:
        if Hm is None:
            Hm, vis_ref = np.asarray(Hb, np.float32), np.asarray(vis)
        else:
            n = min(Hm.shape[0], Hb.shape[0])
            Hm, vis_ref = (Hm[:n] + np.asarray(Hb, np.float32)[:n]).astype(np.float32), vis_ref[:n]
    if Hm is None:
        raise RuntimeError("No keyframes available for POP-GS prior (blocks).")
    return Hm, vis_ref


# ---- map update: get_pointcloud + initialize_new_params (models/SLAM/gaussian.py:75-142, 299-318), numpy ---------

def get_pointcloud(color, depth, intrinsics, w2c, transform_pts=True, downsample=1, mask=None,
                   compute_mean_sq_dist=False):
    """FP32 restatement, operation by operation ("projective" mean_sq_dist).  color [3,H,W], depth [1,H,W]."""
    f = np.float32
    color, depth, K = np.asarray(color, f), np.asarray(depth, f), np.asarray(intrinsics, f)
    H, W = color.shape[1], color.shape[2]
    CX, CY, FX, FY = K[0][2], K[1][2], K[0][0], K[1][1]
    xg, yg = np.meshgrid(np.arange(0, W, downsample).astype(f), np.arange(0, H, downsample).astype(f), indexing="xy")
    xx, yy = ((xg - CX) / FX).reshape(-1), ((yg - CY) / FY).reshape(-1)
    z = depth[0, ::downsample, ::downsample].reshape(-1)
    pts = np.stack((xx * z, yy * z, z), axis=-1).astype(f)
    if transform_pts:
        c2w = np.linalg.inv(np.asarray(w2c, np.float64)).astype(f)
        pts4 = np.concatenate((pts, np.ones((pts.shape[0], 1), f)), axis=1)
        pts = (c2w @ pts4.T).T[:, :3].astype(f)
    msd = None
    if compute_mean_sq_dist:
        sg = (f(downsample) * z / ((FX + FY) / f(2))).astype(f)
        msd = (sg * sg).astype(f)
    cols = np.transpose(color[:, ::downsample, ::downsample], (1, 2, 0)).reshape(-1, 3)
    cld = np.concatenate((pts, cols), axis=-1).astype(f)
    if mask is not None:
        m = np.asarray(mask).reshape(H, W).astype(bool)
        Hd, Wd = H // downsample, W // downsample
        m = m[:Hd * downsample, :Wd * downsample].reshape(Hd, downsample, Wd, downsample).any(axis=(1, 3)).reshape(-1)
        if m.sum() > 0:
            cld = cld[m]
            if msd is not None:
                msd = msd[m]
    return (cld, msd) if compute_mean_sq_dist else cld


def initialize_new_params(new_pt_cld, mean3_sq_dist, isotropic=False):
    n = new_pt_cld.shape[0]
    ls = np.log(np.sqrt(np.asarray(mean3_sq_dist, np.float32))).astype(np.float32)
    return {"means3D": new_pt_cld[:, :3], "rgb_colors": new_pt_cld[:, 3:6],
            "unnorm_rotations": np.tile(np.array([1, 0, 0, 0], np.float32), (n, 1)),
            "logit_opacities": np.zeros((n, 1), np.float32),
            "log_scales": np.tile(ls[:, None], (1, 1 if isotropic else 3))}


# ---- occupancy update and masked visibility counter (numpy restatements, TEST INFRASTRUCTURE) ----
def _r32(x):
    return np.asarray(x, dtype=np.float32)


def _affine_f32(m, x, y, z):
    """fma(m2, z, fma(m1, y, m0 * x)) + m3 in float32: the k-sequential order the CUDA path fixes for `c2w @ p`
    (float64 evaluation of each fma, rounded once, is the correctly rounded float32 fma for float32 inputs)."""
    m = np.asarray(m, dtype=np.float32)
    t = _r32(np.float64(m[0]) * np.float64(x))
    t = _r32(np.float64(m[1]) * np.float64(y) + np.float64(t))
    t = _r32(np.float64(m[2]) * np.float64(z) + np.float64(t))
    return _r32(t + m[3])


def cv_line_8(x1, y1, x2, y2):
    """Pixels of cv2.line(img, (x1, y1), (x2, y2), color, thickness=1, lineType=LINE_8): OpenCV's LineIterator, 8-connected,
    leftToRight (imgproc/src/drawing.cpp: Line() -> LineIterator(img, pt1, pt2, 8, true)).  OpenCV is not in this image:
    parity of this part is UNPINNED; the algorithm is restated from the published source."""
    if x2 < x1:
        x1, y1, x2, y2 = x2, y2, x1, y1
    dx, dy = x2 - x1, y2 - y1
    ystep = -1 if dy < 0 else 1
    dy = abs(dy)
    steep = dy > dx
    major, minor = (dy, dx) if steep else (dx, dy)
    err, plus, minus = major - 2 * minor, 2 * major, -2 * minor
    x, y, out = x1, y1, []
    for _ in range(major + 1):
        out.append((x, y))
        m = err < 0
        err += minus + (plus if m else 0)
        if steep:
            y += ystep
            x += 1 if m else 0
        else:
            x += 1
            y += ystep if m else 0
    return out


def update_occ_map(occ_map, depth, c2w, intr, cell_size, map_center, height_lower, height_upper, far, planes, downsample=1):
    """AstarPlanner.update_occ_map (planning/astar.py:202-303) on numpy arrays, float32 operation by operation.
    occ_map [3,gz,gx] is updated in place; returns the camera cell (x, z)."""
    occ_map_in = occ_map
    gz, gx = occ_map.shape[1], occ_map.shape[2]
    c2w = _r32(c2w); mc = _r32(map_center); cell = np.float32(cell_size)
    fx, fy, cx, cy = (np.float32(v) for v in intr)
    cam_x = int(_r32(_r32(_r32(c2w[0, 3] - mc[0]) / cell) + np.float32(gx // 2)))
    cam_z = int(_r32(_r32(_r32(c2w[2, 3] - mc[1]) / cell) + np.float32(gz // 2)))
    occ_map_in[2, cam_z - 1:cam_z + 2, cam_x - 1:cam_x + 2] = 1e3                      # :214
    d = _r32(depth).reshape(depth.shape[-2], depth.shape[-1])[::downsample, ::downsample]
    ys, xs = np.meshgrid(np.arange(0, depth.shape[-2], downsample, dtype=np.float32),
                         np.arange(0, depth.shape[-1], downsample, dtype=np.float32), indexing="ij")
    xx, yy = _r32((xs - cx) / fx), _r32((ys - cy) / fy)
    cnt = np.zeros((2, gz, gx), np.int64)                                               # [free, occupied]
    for k in range(11):
        dz = _r32(np.float32(planes[k]) * d)
        m = (dz > 0) & (dz < np.float32(far))
        x, y = _r32(xx * dz), _r32(yy * dz)
        wx = _affine_f32(c2w[0], x, y, dz); wy = _affine_f32(c2w[1], x, y, dz); wz = _affine_f32(c2w[2], x, y, dz)
        m &= (wy >= np.float32(height_lower)) & (wy <= np.float32(height_upper))
        bx = _r32(np.floor(_r32(_r32(wx - mc[0]) / cell)) + np.float32((gx - 1) / 2.0)).astype(np.int32).clip(0, gx - 1)
        bz = _r32(np.floor(_r32(_r32(wz - mc[1]) / cell)) + np.float32((gz - 1) / 2.0)).astype(np.int32).clip(0, gz - 1)
        np.add.at(cnt[0 if k < 10 else 1], (bz[m], bx[m]), 1)
    tmp = np.zeros((3, gz, gx), np.float32)
    tmp[2] = np.where(cnt[0] > 0, _r32(np.float32(0.01) * _r32(cnt[0].astype(np.float32) + np.float32(1e-5))), np.float32(0))
    tmp[1] = np.where(cnt[1] > 0, _r32(_r32(cnt[1].astype(np.float32) + np.float32(1e-5)) * np.float32(100)), np.float32(0))
    canvas = np.zeros((gz, gx), bool)
    cxx, czz = min(gx - 1, max(0, cam_x)), min(gz - 1, max(0, cam_z))
    for z, x in zip(*np.nonzero(cnt[1])):
        for (px, pz) in cv_line_8(int(x), int(z), cxx, czz):
            canvas[pz, px] = True
    tmp[2][canvas] = 1.0
    den = _r32(_r32(_r32(tmp[0] + tmp[1]) + tmp[2]) + np.float32(1e-5))
    occ_map_in += _r32(tmp / den)
    return cam_x, cam_z


def count_gaussians_vs_mask(means3D, logit_opacities, w2c, intr, mask, opacity_thresh=0.01):
    """models/SLAM/gaussian_object.py:359-428 on numpy arrays (float32, one rounding per torch op).
    Returns dict(in, out, active, idx_in, idx_out) with the reference's index order."""
    H, W = mask.shape[-2], mask.shape[-1]
    lo = _r32(logit_opacities).reshape(-1)
    alpha = _r32(np.float32(1) / _r32(np.float32(1) + np.exp(-lo, dtype=np.float32)))
    active = alpha >= np.float32(opacity_thresh)
    idx_active = np.nonzero(active)[0]
    X = _r32(means3D)[active]
    w2c = _r32(w2c)
    xc = _affine_f32(w2c[0], X[:, 0], X[:, 1], X[:, 2]); yc = _affine_f32(w2c[1], X[:, 0], X[:, 1], X[:, 2])
    zc = _affine_f32(w2c[2], X[:, 0], X[:, 1], X[:, 2])
    zz = np.maximum(zc, np.float32(1e-6))
    fx, fy, cx, cy = (np.float32(v) for v in intr)
    u = _r32(_r32(fx * _r32(xc / zz)) + cx); v = _r32(_r32(fy * _r32(yc / zz)) + cy)
    ok = (zc > 0) & (u >= 0) & (u < W) & (v >= 0) & (v < H)
    iu = np.rint(u[ok]).astype(np.int64).clip(0, W - 1); iv = np.rint(v[ok]).astype(np.int64).clip(0, H - 1)
    inm = np.asarray(mask, bool).reshape(H, W)[iv, iu]
    idx_ok = idx_active[ok]
    idx_in = idx_ok[inm]
    idx_out = np.concatenate([idx_ok[~inm], idx_active[~ok]])
    return {"in": int(idx_in.size), "out": int(idx_out.size), "active": int(active.sum()), "idx_in": idx_in, "idx_out": idx_out}
