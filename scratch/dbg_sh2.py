import numpy as np, torch, sys, os
sys.path.insert(0, os.getcwd())
import fisher_nerf_customized_b200 as fb
from tests import common
from tests.test_gpu_api_options import _fwd, _settings, _t
from oracle import build_ref
ref = build_ref.import_ref("ref_dgr")
W=H=64; P=4000
sc = common.random_view_scene(P, 40, W, H)
t = _t(torch, sc)
rs = _settings(fb, torch, W, H, degree=0)
for M in (1, 16):
    shs = torch.from_numpy(np.random.default_rng(0).normal(0,0.5,size=(P,M,3)).astype(np.float32)).cuda()
    e = torch.Tensor([])
    for name, campos in (("cuda campos", rs.campos), ("cpu campos", rs.campos.cpu())):
        try:
            b = ref.rasterize_gaussians(rs.bg, t["means3D"], e, t["opacities"], t["scales"], t["rotations"], 1.0, e,
                 rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, H, W, shs, 0, campos, False)
            torch.cuda.synchronize()
            print(M, name, "img max", b[1].abs().max().item(), "R", b[0])
        except Exception as ex:
            print(M, name, "EXC", ex)
    print("campos", rs.campos, rs.campos.dtype, rs.campos.is_contiguous(), "bg", rs.bg)
    # all-positive sh
    shp = shs.abs() + 1.0
    b = ref.rasterize_gaussians(rs.bg, t["means3D"], e, t["opacities"], t["scales"], t["rotations"], 1.0, e,
         rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, H, W, shp, 0, rs.campos, False)
    print(M, "positive sh: img max", b[1].abs().max().item())
