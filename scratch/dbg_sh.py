import numpy as np, torch, sys, os
sys.path.insert(0, os.getcwd())
import fisher_nerf_customized_b200 as fb
from tests import common
from tests.test_gpu_api_options import _fwd, _settings, _t
from oracle import build_ref
ref = build_ref.import_ref("ref_dgr")
W=H=64; P=4000; M=16
sc = common.random_view_scene(P, 40, W, H)
shs = torch.from_numpy(np.random.default_rng(0).normal(0,0.5,size=(P,M,3)).astype(np.float32)).cuda()
for deg in (0,3):
    rs = _settings(fb, torch, W, H, degree=deg)
    t = _t(torch, sc)
    a = _fwd(fb._C, torch, rs, t, shs=shs, degree=deg)
    b = _fwd(ref, torch, rs, t, shs=shs, degree=deg)
    torch.cuda.synchronize()
    print(deg, "ours", a[1].abs().max().item(), a[1].mean().item(), "ref", b[1].abs().max().item(), b[1].mean().item(), "R", a[0], b[0])
    # ref with colours computed by us for comparison
    e = torch.Tensor([])
    print(" e.data_ptr", e.data_ptr(), "cuda empty", torch.empty(0, device="cuda").data_ptr())
    b2 = ref.rasterize_gaussians(rs.bg, t["means3D"], torch.empty(0, device="cuda"), t["opacities"], t["scales"], t["rotations"], 1.0, torch.empty(0, device="cuda"),
         rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, H, W, shs, deg, rs.campos, False)
    print(" ref with cuda-empty colours:", b2[1].abs().max().item(), b2[1].mean().item())
print("---- geom dump")
rs = _settings(fb, torch, W, H, degree=0)
b = _fwd(ref, torch, rs, t, shs=shs, degree=0)
geom = b[3]
print("geom bytes", geom.numel())
rgb = geom[236032:236032+48000].view(torch.float32)
print("ref rgb stats", rgb.abs().max().item(), rgb.mean().item(), torch.isnan(rgb).sum().item())
cl = geom[16000:28000]
print("clamped sum", cl.sum().item())
img = b[5]
nc = img[H*W*4:H*W*8].view(torch.int32)
print("n_contrib max", nc.max().item(), "final_T min", img[:H*W*4].view(torch.float32).min().item())
print("depth img", b[6].abs().max().item())
