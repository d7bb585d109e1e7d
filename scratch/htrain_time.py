import numpy as np, torch, sys, os, time
sys.path.insert(0, os.getcwd())
import fisher_nerf_customized_b200 as fb
from fisher_nerf_customized_b200 import synth
P=500000
sc = synth.activate(synth.make_scene(P))
t = {k: torch.from_numpy(v).cuda() for k, v in sc.items()}
gmap = fb.GaussianMap(t["means3D"], t["rgb_colors"], t["rotations"], t["opacities"], t["scales"])
rs = fb.setup_camera(256,256, synth.intrinsics(256,256), np.eye(4, dtype=np.float32))
kf = torch.from_numpy(synth.w2c_from_c2w(synth.make_poses(32, seed=synth.SEED_POSES+1))).cuda()
for V in (1, 8, 16, 32):
    s = fb.FisherScorer(rs, C_H=4, views_per_chunk=V)
    H = s.compute_H_train(gmap, kf)
    torch.cuda.synchronize(); t0=time.perf_counter()
    for _ in range(3): H = s.compute_H_train(gmap, kf)
    torch.cuda.synchronize(); dt=(time.perf_counter()-t0)/3
    print(f"V={V}: {dt*1e3:.1f} ms for 32 keyframes = {32/dt:.0f} keyframes/s; sum(H)={float(H.double().sum()):.6e}")
from fisher_nerf_customized_b200 import _lib
s = fb.FisherScorer(rs, C_H=4, views_per_chunk=32)
s.compute_H_train(gmap, kf)
_lib.profile_enable(True); _lib.profile_read()
s.compute_H_train(gmap, kf)
prof = _lib.profile_read(); _lib.profile_enable(False)
for k,v in prof.items():
    if v[1]: print(f"  {k:18s} {v[0]*1e3/v[1]:9.1f} us/launch x{v[1]}")
