"""Lane efficiency of k_tile_score1's pair loop on the bench workload: run with FISHER_B200_LIB pointing at a
-DFB_SCORE_STATS={1,2,3} build; prints the per-view count the build returns instead of scores."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import fisher_nerf_customized_b200 as fb
from fisher_nerf_customized_b200 import synth
W = H = 256
sc = synth.activate(synth.make_scene(500_000))
t = {k: torch.from_numpy(v).cuda() for k, v in sc.items()}
gmap = fb.GaussianMap(t["means3D"], t["rgb_colors"], t["rotations"], t["opacities"], t["scales"])
rs = fb.setup_camera(W, H, synth.intrinsics(W, H), np.eye(4, dtype=np.float32))
scorer = fb.FisherScorer(rs, C_H=4, views_per_chunk=64)
w2c = torch.from_numpy(synth.w2c_from_c2w(synth.make_poses(64, seed=5678))).cuda()
Hinv = torch.ones((500_000, 4), device="cuda")
s = scorer.score_candidates(gmap, w2c, Hinv).double()
print(os.environ.get("FISHER_B200_LIB", "default").split("_")[-1], "per view:", float(s.mean()), "instances per view", scorer.last_instances / 64)
