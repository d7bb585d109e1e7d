import numpy as np, torch, sys, os
sys.path.insert(0, os.getcwd())
import fisher_nerf_customized_b200 as fb
from fisher_nerf_customized_b200 import synth
P=500000
sc = synth.activate(synth.make_scene(P))
t = {k: torch.from_numpy(v).cuda() for k, v in sc.items()}
gmap = fb.GaussianMap(t["means3D"], t["rgb_colors"], t["rotations"], t["opacities"], t["scales"])
rs = fb.setup_camera(256,256, synth.intrinsics(256,256), np.eye(4, dtype=np.float32))
w2c = torch.from_numpy(synth.w2c_from_c2w(synth.make_poses(256, seed=synth.SEED_POSES))).cuda()
Hinv = torch.ones((P,4), device="cuda")
mx=[]; tot=[]
for k in range(0,256,1):
    s = fb.FisherScorer(rs, C_H=4, views_per_chunk=1)
    s.score_candidates(gmap, w2c[k:k+1], Hinv)
    mx.append(s.last_status[2]); tot.append(s.last_status[3])
mx=np.array(mx); tot=np.array(tot)
print("per-view max tile list: median", np.median(mx), "p90", np.percentile(mx,90), "max", mx.max())
print("per-view instances: median", np.median(tot), "p90", np.percentile(tot,90), "max", tot.max())
print("views with max tile > 4096:", (mx>4096).sum(), " > 8192:", (mx>8192).sum())
