"""Shared test inputs: seeded small scenes in the reference's parameterisation."""
import numpy as np

from fisher_nerf_customized_b200 import synth


def small_scene(P, seed=0, box=(6.0, 3.0, 6.0), focal=32.0):
    raw = synth.make_scene(P, seed=seed, box=box, focal=focal, mean_depth=2.0)
    return synth.activate(raw)


def random_view_scene(P, seed=0, W=64, H=64):
    """Gaussians in front of an identity camera (camera frame == world frame), all kinds of sizes,
    some behind the camera and some outside the frustum: exercises every cull path of preprocess."""
    rng = np.random.default_rng(seed)
    means = np.c_[rng.uniform(-3, 3, P), rng.uniform(-3, 3, P), rng.uniform(-1.0, 6.0, P)].astype(np.float32)
    q = rng.normal(size=(P, 4)).astype(np.float32)
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    scales = np.exp(rng.uniform(np.log(0.01), np.log(0.4), size=(P, 3))).astype(np.float32)
    op = (1 / (1 + np.exp(-rng.normal(0, 1.5, size=(P, 1))))).astype(np.float32)
    col = rng.uniform(0, 1, size=(P, 3)).astype(np.float32)
    return dict(means3D=means, rgb_colors=col, rotations=q, opacities=op, scales=scales)


def K_for(W, H):
    return synth.intrinsics(W, H)


def look_at_w2c(eye, target, down=(0.0, 1.0, 0.0)):
    eye, target, down = (np.asarray(a, dtype=np.float64) for a in (eye, target, down))
    f = target - eye
    f /= np.linalg.norm(f)
    r = np.cross(down, f)
    r /= np.linalg.norm(r)
    d = np.cross(f, r)
    c2w = np.eye(4)
    c2w[:3, 0], c2w[:3, 1], c2w[:3, 2], c2w[:3, 3] = r, d, f, eye
    return np.linalg.inv(c2w).astype(np.float32)
