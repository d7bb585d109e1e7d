"""Where does a per-rank share of a strong-scaled planning step go?  One GPU, 4 keyframes + 32 candidates (what a rank
of 8 gets of BASELINE cfg 3): CUDA-event time of ShardedScorer.pose_eval against the sum of its kernels' own times."""
import sys, os, time, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import fisher_nerf_customized_b200 as fb
from fisher_nerf_customized_b200 import synth, _lib
P = 500_000
sc = synth.activate(synth.make_scene(P))
t = {k: torch.from_numpy(v).cuda() for k, v in sc.items()}
gmap = fb.GaussianMap(t["means3D"], t["rgb_colors"], t["rotations"], t["opacities"], t["scales"])
rs = fb.setup_camera(256, 256, synth.intrinsics(256, 256), np.eye(4, dtype=np.float32))
scorer = fb.FisherScorer(rs, C_H=4, views_per_chunk=64, chunk_lanes=int(os.environ.get("FB_CHUNK_LANES", "2")))
print("chunk_lanes", scorer.chunk_lanes)
sh = fb.ShardedScorer(scorer)
kf_all = torch.from_numpy(synth.w2c_from_c2w(synth.make_poses(32, seed=synth.SEED_POSES + 1))).cuda()
c2w_all = torch.from_numpy(synth.make_poses(256, seed=synth.SEED_POSES)).cuda()
for nkf, K in ((32, 256), (4, 32), (1, 8)):
    kf, c2w = kf_all[:nkf], c2w_all[:K]
    for _ in range(3):
        sh.pose_eval(gmap, c2w, kf, 0.1)
    torch.cuda.synchronize()
    n = 20
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter(); a.record()
    for _ in range(n):
        sh.pose_eval(gmap, c2w, kf, 0.1)
    b.record(); b.synchronize()
    wall = (time.perf_counter() - t0) / n * 1e3
    ev = a.elapsed_time(b) / n
    _lib.profile_enable(True); _lib.profile_read()
    for _ in range(n):
        sh.pose_eval(gmap, c2w, kf, 0.1)
    prof = _lib.profile_read(); _lib.profile_enable(False)
    ksum = sum(v[0] for v in prof.values()) / n
    # host-only cost: the same calls with the GPU idle in between (sync after each)
    print(f"kf {nkf:3d} cand {K:4d}: wall {wall:.3f} ms, events {ev:.3f} ms, kernels alone {ksum:.3f} ms "
          f"({ {k: round(v[0] / n * 1e3) for k, v in prof.items()} } us)", flush=True)
