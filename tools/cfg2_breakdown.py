"""Per-kernel CUDA-event times of the drop-in forward + backward (BASELINE cfg 2: 200k Gaussians, 256x256), power 1 and 2."""
import sys, os, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import fisher_nerf_customized_b200 as fb
from fisher_nerf_customized_b200 import synth, _lib
P=200000; W=H=256
sc = synth.activate(synth.make_scene(P))
t = {k: torch.from_numpy(v).cuda() for k, v in sc.items()}
rs = fb.setup_camera(W, H, synth.intrinsics(W, H), np.eye(4, dtype=np.float32))
for power in (1, 2):
    def run():
        rv = dict(means3D=t["means3D"].clone().requires_grad_(True), colors_precomp=t["rgb_colors"].clone().requires_grad_(True),
                  rotations=t["rotations"].clone().requires_grad_(True), opacities=t["opacities"].clone().requires_grad_(True),
                  scales=t["scales"].clone().requires_grad_(True), means2D=torch.zeros_like(t["means3D"], requires_grad=True))
        im, radius, _ = fb.GaussianRasterizer(raster_settings=rs, backward_power=power)(**rv)
        im.backward(gradient=torch.ones_like(im) * 1e-3)
    for _ in range(3): run()
    _lib.profile_enable(True); _lib.profile_read()
    for _ in range(5): run()
    prof = _lib.profile_read(); _lib.profile_enable(False)
    print("power", power, {k: round(v[0] / 5 * 1e3) for k, v in prof.items() if v[1]})
