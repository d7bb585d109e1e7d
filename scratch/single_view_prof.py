import numpy as np, torch, sys, os, time
sys.path.insert(0, os.getcwd())
import fisher_nerf_customized_b200 as fb
from fisher_nerf_customized_b200 import synth, _lib
P, W, H = 200_000, 256, 256
sc = synth.activate(synth.make_scene(P))
t = {k: torch.from_numpy(v).cuda() for k, v in sc.items()}
w2c = torch.from_numpy(synth.w2c_from_c2w(synth.make_poses(1))[0]).cuda()
pts4 = torch.cat((t["means3D"], torch.ones(P, 1, device="cuda")), dim=1)
tp = (w2c @ pts4.T).T[:, :3].contiguous()
rs = fb.setup_camera(W, H, synth.intrinsics(W, H), np.eye(4, dtype=np.float32))
e = torch.Tensor([]); dpix = torch.ones((3, H, W), device="cuda") * 1e-3
def run(power):
    R, color, radii, geom, binning, img, depth = fb._C.rasterize_gaussians(rs.bg, tp, t["rgb_colors"], t["opacities"], t["scales"], t["rotations"], 1.0, e, rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, H, W, e, 0, rs.campos, False)
    g = fb._C.rasterize_gaussians_backward(rs.bg, tp, radii, t["rgb_colors"], t["scales"], t["rotations"], 1.0, e, rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, dpix, e, 0, rs.campos, geom, R, binning, img, power)
    return R
for _ in range(5): R = run(2)
torch.cuda.synchronize(); t0 = time.perf_counter()
for _ in range(50): run(2)
torch.cuda.synchronize(); print("fwd+bwd wall ms", (time.perf_counter()-t0)/50*1e3, "R", R)
_lib.profile_enable(True); _lib.profile_read()
for _ in range(10): run(2)
prof = _lib.profile_read(); _lib.profile_enable(False)
for k,v in prof.items():
    if v[1]: print(f"  {k:18s} {v[0]*1e3/v[1]:9.1f} us/launch x{v[1]}")
