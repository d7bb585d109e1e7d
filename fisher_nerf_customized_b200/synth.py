"""Seeded synthetic Gaussian maps, candidate poses and point clouds of the shapes BASELINE.json names
(SURVEY.md 8(d)).  numpy only; used by bench.py, the tests and the golden-vector script so that every
arm sees identical inputs.

Distributions mirror the reference where it defines them:
  scales        projective initialisation log(sqrt(mean_sq_dist)), depth/focal (models/SLAM/gaussian.py:110-117,161)
  candidates    generate_candidate: ring of yaw-only poses around free-space centres at fixed camera
                height, looking at the centre (planning/astar.py:1383-1425)
  camera        fx = fy = cx = cy = W/2 for W = H = 256 (configs/base_config.py:179-190)
"""
import numpy as np

SEED_SCENE, SEED_POSES, SEED_POINTS = 1234, 5678, 91011


def intrinsics(W=256, H=256):
    return np.array([[W / 2.0, 0.0, W / 2.0], [0.0, H / 2.0, H / 2.0], [0.0, 0.0, 1.0]], dtype=np.float32)


def _shell_points(rng, n, box, frac_shell=0.7, jitter=0.02, n_furniture=40):
    """70 % on the six faces of the room box (normal jitter 2 cm), 30 % on the faces of random interior
    axis-aligned boxes: an indoor-like shell, so depth complexity is realistic rather than uniform fog."""
    bx, by, bz = box
    half = np.array([bx / 2, by / 2, bz / 2])

    def on_box_faces(m, centre, h):
        # pick a face with probability proportional to its area
        areas = np.array([h[1] * h[2], h[1] * h[2], h[0] * h[2], h[0] * h[2], h[0] * h[1], h[0] * h[1]])
        face = rng.choice(6, size=m, p=areas / areas.sum())
        p = rng.uniform(-1, 1, size=(m, 3)) * h
        axis = face // 2
        sign = np.where(face % 2 == 0, -1.0, 1.0)
        p[np.arange(m), axis] = sign * h[axis] + rng.normal(0, jitter, size=m)
        return p + centre

    n_shell = int(round(n * frac_shell))
    pts = [on_box_faces(n_shell, np.zeros(3), half)]
    n_f = n - n_shell
    if n_f > 0:
        per = np.full(n_furniture, n_f // n_furniture)
        per[: n_f - per.sum()] += 1
        for m in per:
            if m == 0:
                continue
            h = rng.uniform(0.3, 1.5, size=3) / 2
            h[1] = min(h[1], half[1])
            c = rng.uniform(-1, 1, size=3) * (half - h)
            pts.append(on_box_faces(int(m), c, h))
    return np.concatenate(pts, axis=0)


def make_scene(P, seed=SEED_SCENE, kind="shell", box=(10.0, 3.0, 10.0), focal=128.0, mean_depth=3.0):
    """Raw map parameters in the reference's checkpoint key set (gaussian.py:156-168)."""
    rng = np.random.default_rng(seed)
    if kind == "shell":
        means = _shell_points(rng, P, box)
    elif kind == "fog":
        means = rng.uniform(-0.5, 0.5, size=(P, 3)) * np.array(box)
    else:
        raise ValueError(kind)
    perm = rng.permutation(P)           # no spatial order in the index (as after densification)
    means = means[perm]
    base = rng.uniform(0.5, 2.0, size=(P, 1)) * mean_depth / focal
    aniso = rng.uniform(0.5, 2.0, size=(P, 3))
    q = rng.normal(size=(P, 4))
    return dict(means3D=means.astype(np.float32),
                rgb_colors=rng.uniform(0, 1, size=(P, 3)).astype(np.float32),
                unnorm_rotations=q.astype(np.float32),
                logit_opacities=rng.normal(0, 1.5, size=(P, 1)).astype(np.float32),
                log_scales=np.log(base * aniso).astype(np.float32))


def activate(params):
    """The activated form compute_Hessian feeds the rasteriser (gaussian.py:1529-1533), float32."""
    q = params["unnorm_rotations"].astype(np.float32)
    n = np.maximum(np.sqrt((q * q).sum(1, keepdims=True, dtype=np.float32)), np.float32(1e-12))
    scales = np.exp(params["log_scales"].astype(np.float32))
    if scales.shape[-1] == 1:
        scales = np.tile(scales, (1, 3))
    return dict(means3D=params["means3D"].astype(np.float32), rgb_colors=params["rgb_colors"].astype(np.float32),
                rotations=(q / n).astype(np.float32),
                opacities=(1.0 / (1.0 + np.exp(-params["logit_opacities"].astype(np.float32)))).astype(np.float32),
                scales=scales.astype(np.float32))


def make_poses(K, seed=SEED_POSES, box=(10.0, 3.0, 10.0), n_centres=8):
    """c2w [K,4,4] float32: camera x right, y down, z forward; yaw-only, camera height 0 (box mid-height);
    positions on rings of radius U[0.5,1]*expansion around random free-space centres, looking at the
    centre (yaw = theta + pi), clamped to stay inside the room."""
    rng = np.random.default_rng(seed)
    bx, _, bz = box
    centres = rng.uniform(-0.5, 0.5, size=(n_centres, 2)) * np.array([bx - 3.0, bz - 3.0])
    out = np.zeros((K, 4, 4), np.float32)
    for k in range(K):
        c = centres[k % n_centres]
        theta = rng.uniform(0, 2 * np.pi)
        r = rng.uniform(0.5, 1.0) * 1.2
        pos = np.array([c[0] + r * np.sin(theta), 0.0, c[1] + r * np.cos(theta)])
        pos[0] = np.clip(pos[0], -bx / 2 + 0.3, bx / 2 - 0.3)
        pos[2] = np.clip(pos[2], -bz / 2 + 0.3, bz / 2 - 0.3)
        yaw = theta + np.pi
        fwd = np.array([np.sin(yaw), 0.0, np.cos(yaw)])
        down = np.array([0.0, 1.0, 0.0])
        right = np.cross(down, fwd)
        m = np.eye(4)
        m[:3, 0], m[:3, 1], m[:3, 2], m[:3, 3] = right, down, fwd, pos
        out[k] = m.astype(np.float32)
    return out


def w2c_from_c2w(c2w):
    """float32 inverse, as torch.linalg.inv on float32 poses does in pose_eval (gaussian.py:1364)."""
    return np.stack([np.linalg.inv(m.astype(np.float32)) for m in np.asarray(c2w)]).astype(np.float32)


def make_points(N, seed=SEED_POINTS, box=(10.0, 3.0, 10.0)):
    rng = np.random.default_rng(seed)
    return _shell_points(rng, N, box)[rng.permutation(N)].astype(np.float32)
