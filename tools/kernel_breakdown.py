"""Per-kernel CUDA-event times of one scoring call: python tools/kernel_breakdown.py P K C_H [box_x box_z]
(FB_BREAKDOWN_HTRAIN=1: of one compute_H_train call over K keyframes instead)"""
import sys, os, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import fisher_nerf_customized_b200 as fb
from fisher_nerf_customized_b200 import synth, _lib
P, K, C_H = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
box = (float(sys.argv[4]), 3.0, float(sys.argv[5])) if len(sys.argv) > 5 else (10.0, 3.0, 10.0)
sc = synth.activate(synth.make_scene(P, box=box))
t = {k: torch.from_numpy(v).cuda() for k, v in sc.items()}
gmap = fb.GaussianMap(t["means3D"], t["rgb_colors"], t["rotations"], t["opacities"], t["scales"])
rs = fb.setup_camera(256, 256, synth.intrinsics(256, 256), np.eye(4, dtype=np.float32))
scorer = fb.FisherScorer(rs, C_H=C_H, views_per_chunk=64)
w2c = torch.from_numpy(synth.w2c_from_c2w(synth.make_poses(K, box=box))).cuda()
H_inv = torch.rand((P, C_H), device="cuda") + 0.1
HT = os.environ.get("FB_BREAKDOWN_HTRAIN") == "1"
run = (lambda: scorer.compute_H_train(gmap, w2c)) if HT else (lambda: scorer.score_candidates(gmap, w2c, H_inv))
for _ in range(2):
    run()
_lib.profile_enable(True); _lib.profile_read()
run()
prof = _lib.profile_read(); _lib.profile_enable(False)
tot = sum(v[0] for v in prof.values())
print(f"P {P} K {K} C_H {C_H}: {tot:.2f} ms = {K / tot * 1e3:.0f} views/s; instances/view {scorer.last_instances / K:.0f}, visible/view "
      f"{scorer.last_visible / K:.0f}, vis_capacity {scorer.vis_capacity}, us per launch:",
      {k: round(v[0] / v[1] * 1e3) for k, v in prof.items()}, flush=True)
