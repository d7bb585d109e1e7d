"""Instance-count statistics of the synthetic cfg-3 scene at different binning granularities (numpy, CPU).
Design aid for the scoring path: how many (Gaussian, bin) instances a view has when bins are the reference's
16x16 tiles with its 3-sigma rectangle, or BXxBY pixel blocks with the exact alpha >= 1/255 footprint box."""
import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fisher_nerf_customized_b200 import synth

P = int(sys.argv[1]) if len(sys.argv) > 1 else 500_000
NV = int(sys.argv[2]) if len(sys.argv) > 2 else 4
W = H = 256
fx = fy = 128.0
sc = synth.activate(synth.make_scene(P))
poses = synth.w2c_from_c2w(synth.make_poses(NV, seed=synth.SEED_POSES))

def quat_R(q):
    r, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    R = np.empty((len(q), 3, 3))
    R[:, 0, 0] = 1 - 2 * (y * y + z * z); R[:, 0, 1] = 2 * (x * y - r * z); R[:, 0, 2] = 2 * (x * z + r * y)
    R[:, 1, 0] = 2 * (x * y + r * z); R[:, 1, 1] = 1 - 2 * (x * x + z * z); R[:, 1, 2] = 2 * (y * z - r * x)
    R[:, 2, 0] = 2 * (x * z - r * y); R[:, 2, 1] = 2 * (y * z + r * x); R[:, 2, 2] = 1 - 2 * (x * x + y * y)
    return R

Rm = quat_R(sc["rotations"].astype(np.float64))
S = sc["scales"].astype(np.float64)
M = Rm * S[:, None, :]
Sigma = M @ M.transpose(0, 2, 1)      # world-frame cov (reference keeps world rotations with camera-frame means)
op = sc["opacities"].reshape(-1).astype(np.float64)

for v in range(NV):
    w2c = poses[v].astype(np.float64)
    pc = sc["means3D"].astype(np.float64) @ w2c[:3, :3].T + w2c[:3, 3]
    z = pc[:, 2]
    ok = z > 0.001
    tx = np.clip(pc[:, 0] / z, -1.3, 1.3) * z; ty = np.clip(pc[:, 1] / z, -1.3, 1.3) * z
    J = np.zeros((P, 2, 3))
    J[:, 0, 0] = fx / z; J[:, 0, 2] = -fx * tx / (z * z); J[:, 1, 1] = fy / z; J[:, 1, 2] = -fy * ty / (z * z)
    cov = J @ Sigma @ J.transpose(0, 2, 1)    # view = identity in the scoring path
    a = cov[:, 0, 0] + 0.3; b = cov[:, 0, 1]; c = cov[:, 1, 1] + 0.3
    det = a * c - b * b
    mid = 0.5 * (a + c)
    lam = mid + np.sqrt(np.maximum(0.1, mid * mid - det))
    rad = np.ceil(3 * np.sqrt(lam))
    px = fx * pc[:, 0] / z + W / 2 - 0.5; py = fy * pc[:, 1] / z + H / 2 - 0.5
    x0 = np.clip(((px - rad) / 16).astype(int), 0, 16); x1 = np.clip(((px + rad + 15) / 16).astype(int), 0, 16)
    y0 = np.clip(((py - rad) / 16).astype(int), 0, 16); y1 = np.clip(((py + rad + 15) / 16).astype(int), 0, 16)
    vis = ok & (x1 > x0) & (y1 > y0)
    R_ref = ((x1 - x0) * (y1 - y0))[vis].sum()
    ca, cb, cc = c / det, -b / det, a / det
    tau2 = 2 * np.log(np.maximum(255 * op, 1.0)) + 1e-3
    dq = ca * cc - cb * cb
    hx = np.sqrt(tau2 * cc / dq) * 1.001 + 0.05; hy = np.sqrt(tau2 * ca / dq) * 1.001 + 0.05
    live = vis & (op >= 1 / 255)
    # footprint box clipped to the 3-sigma tile rectangle and the image, in pixel-centre coordinates
    bx0 = np.maximum(np.ceil(px - hx), x0 * 16); bx1 = np.minimum(np.floor(px + hx), np.minimum(x1 * 16, W) - 1)
    by0 = np.maximum(np.ceil(py - hy), y0 * 16); by1 = np.minimum(np.floor(py + hy), np.minimum(y1 * 16, H) - 1)
    live &= (bx1 >= bx0) & (by1 >= by0)
    out = [f"view {v}: visible {vis.sum()}, R_ref(16x16, 3 sigma) {R_ref}, live {live.sum()}"]
    for (BX, BY) in [(16, 16), (8, 8), (8, 4), (4, 4)]:
        nx = (bx1 // BX - bx0 // BX + 1)[live]; ny = (by1 // BY - by0 // BY + 1)[live]
        n = (nx * ny)
        # per-bin list lengths
        cnt = np.zeros((H // BY, W // BX), np.int64)
        idx = np.nonzero(live)[0]
        X0 = (bx0[idx] // BX).astype(int); Y0 = (by0[idx] // BY).astype(int)
        nxi = nx.astype(int); nyi = ny.astype(int)
        for dxi in range(int(nxi.max())):
            for dyi in range(int(nyi.max())):
                m = (nxi > dxi) & (nyi > dyi)
                np.add.at(cnt, (Y0[m] + dyi, X0[m] + dxi), 1)
        flat = cnt.reshape(-1)
        # warp groups of 4 adjacent 4x4 blocks (or 2): imbalance = sum of max over groups / mean
        msg = f"  {BX}x{BY}: instances {int(n.sum())} ({n.sum()/vis.sum():.2f}/visible) lists mean {flat.mean():.0f} max {flat.max()} lane-visits {int(n.sum())*BX*BY/1e6:.1f}M"
        if (BX, BY) == (4, 4):
            g = cnt.reshape(H // 4, W // 16, 4)           # 4 horizontally adjacent blocks
            msg += f" | quad max-sum {g.max(2).sum()} (x4 = {4*g.max(2).sum()})"
            g2 = cnt.reshape(H // 8, 2, W // 8, 2).transpose(0, 2, 1, 3).reshape(H // 8, W // 8, 4)  # 2x2 blocks = 8x8 px
            msg += f" 2x2 max-sum {g2.max(2).sum()} (x4 = {4*g2.max(2).sum()})"
            g3 = cnt.reshape(H // 4, W // 8, 2)
            msg += f" pair max-sum {g3.max(2).sum()} (x2 = {2*g3.max(2).sum()})"
        out.append(msg)
    print("\n".join(out), flush=True)
