"""One pass over the kernels OFF the headline bench (drop-in single view, H_train, stage 5, block form, densify) for an
ncu capture: python tools/prof_other_paths.py [reps]"""
import numpy as np, torch, sys, os, time
sys.path.insert(0, os.getcwd())
import fisher_nerf_customized_b200 as fb
from fisher_nerf_customized_b200 import synth, _lib
from tests.golden.make_golden import frame_inputs
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 3
W = H = 256
rs = fb.setup_camera(W, H, synth.intrinsics(W, H), np.eye(4, dtype=np.float32))
e = torch.Tensor([])

def timed(name, fn, n=reps):
    fn(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); b.synchronize()
    print(f"{name:44s} {a.elapsed_time(b) / n:9.3f} ms", flush=True)

# cfg 2: single view fwd+bwd, 200k Gaussians
P = 200_000
sc = synth.activate(synth.make_scene(P)); t = {k: torch.from_numpy(v).cuda() for k, v in sc.items()}
w2c = torch.from_numpy(synth.w2c_from_c2w(synth.make_poses(1))[0]).cuda()
tp = (w2c @ torch.cat((t["means3D"], torch.ones(P, 1, device="cuda")), dim=1).T).T[:, :3].contiguous()
dpix = torch.ones((3, H, W), device="cuda") * 1e-3
def single(power):
    R, color, radii, geom, binning, img, depth = fb._C.rasterize_gaussians(rs.bg, tp, t["rgb_colors"], t["opacities"], t["scales"], t["rotations"], 1.0, e, rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, H, W, e, 0, rs.campos, False)
    fb._C.rasterize_gaussians_backward(rs.bg, tp, radii, t["rgb_colors"], t["scales"], t["rotations"], 1.0, e, rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, dpix, e, 0, rs.campos, geom, R, binning, img, power)
timed("single view fwd+bwd power 1, 200k", lambda: single(1))
timed("single view fwd+bwd power 2, 200k", lambda: single(2))
# block form on the same map
gmap2 = fb.GaussianMap(t["means3D"], t["rgb_colors"], t["rotations"], t["opacities"], t["scales"])
scorer11 = fb.FisherScorer(rs, C_H=11)
gen = torch.Generator(device="cuda").manual_seed(1)
Hb, vis = scorer11.estimate_block_JtJ(gmap2, w2c, K=2, generator=gen)
timed("estimate_block_JtJ K=2, 200k", lambda: scorer11.estimate_block_JtJ(gmap2, w2c, K=2, generator=gen))
timed(f"t_opt_blocks, {Hb.shape[0]} blocks of 11x11", lambda: scorer11.t_opt_blocks(Hb, Hb, 1e-6))
timed(f"d_opt_blocks, {Hb.shape[0]} blocks of 11x11", lambda: scorer11.d_opt_blocks(Hb, Hb, 1e-6))
# H_train: 500k Gaussians x 32 keyframes
P = 500_000
sc = synth.activate(synth.make_scene(P)); t5 = {k: torch.from_numpy(v).cuda() for k, v in sc.items()}
gmap = fb.GaussianMap(t5["means3D"], t5["rgb_colors"], t5["rotations"], t5["opacities"], t5["scales"])
kf = torch.from_numpy(synth.w2c_from_c2w(synth.make_poses(32, seed=synth.SEED_POSES + 1))).cuda()
s32 = fb.FisherScorer(rs, C_H=4, views_per_chunk=32)
timed("compute_H_train 500k x 32 keyframes", lambda: s32.compute_H_train(gmap, kf))
# stage 5
pts = torch.from_numpy(synth.make_points(3_000_000, seed=91011)).cuda()
timed("distCUDA2 3M points", lambda: fb.distCUDA2(pts))
pts1 = torch.from_numpy(synth.make_points(1_000_000, seed=91011)).cuda()
poses = synth.make_poses(64, seed=5678).astype(np.float64)
intr = synth.intrinsics(W, H).astype(np.float32)
timed("count_visible 1M x 64 poses", lambda: fb.count_visible_batch(pts1, poses, intr, W))
# densify
color, depth, K, wc, mask = frame_inputs(H, W, 5)
tc, td, tK, tw, tm = [torch.from_numpy(np.ascontiguousarray(x)).cuda() for x in (color, depth, K, wc, mask)]
store = fb.ResidentParams(capacity=200_000)
def dens():
    store.n = 0
    store.append_from_frame(tc, td, tK, tw, mask=tm)
timed("append_from_frame 256x256, 30 % of the pixels", dens)
