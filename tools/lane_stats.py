"""Work counts of the pair sweep on the cfg-3 workload, from the -DFB_S2_STATS=n measurement builds
(tools/variants.sh stats1="-DFB_S2_STATS=1" ...): per view, the mean of scan steps, ring entries, entries after the
exact test, full bodies and contributing pixel pairs.  Run on the GPU box: python tools/lane_stats.py"""
import os, subprocess, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CODE = r'''
import sys, numpy as np, torch
sys.path.insert(0, %r)
import fisher_nerf_customized_b200 as fb
from fisher_nerf_customized_b200 import synth
import os
P = int(os.environ.get('FB_STATS_P', '500000')); box = (30.0, 3.0, 30.0) if P > 1000000 else (10.0, 3.0, 10.0)
sc = synth.activate(synth.make_scene(P, box=box))
t = {k: torch.from_numpy(v).cuda() for k, v in sc.items()}
gmap = fb.GaussianMap(t["means3D"], t["rgb_colors"], t["rotations"], t["opacities"], t["scales"])
rs = fb.setup_camera(256, 256, synth.intrinsics(256, 256), np.eye(4, dtype=np.float32))
scorer = fb.FisherScorer(rs, C_H=4, views_per_chunk=64)
w2c = torch.from_numpy(synth.w2c_from_c2w(synth.make_poses(64, box=box))).cuda()
s = scorer.score_candidates(gmap, w2c, torch.ones((P, 4), device="cuda"))
print(float(s.double().mean()), scorer.last_instances / 64, scorer.last_visible / 64, 'longest tile list', scorer.last_status[2])
''' % ROOT
names = {1: "scan steps (warp)", 2: "ring entries (box survivors)", 3: "entries after the exact test",
         4: "full bodies", 5: "contributing pixel pairs"}
for n in range(1, 6):
    lib = os.path.join(ROOT, "build_variants", f"lib_stats{n}.so")
    if not os.path.exists(lib):
        continue
    env = dict(os.environ, FISHER_B200_LIB=lib)
    out = subprocess.run([sys.executable, "-c", CODE], env=env, capture_output=True, text=True)
    print(f"{names[n]:32s}", out.stdout.strip() or out.stderr[-300:])
