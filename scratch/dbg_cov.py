import numpy as np, torch, sys, os
sys.path.insert(0, os.getcwd())
import fisher_nerf_customized_b200 as fb
from tests import common
from tests.test_gpu_api_options import _fwd, _bwd, _settings, _t, _cov3d_np
from oracle import build_ref
ref = build_ref.import_ref("ref_dgr")
W, H, P = 96, 64, 6000
sc = common.random_view_scene(P, 50, W, H)
t = _t(torch, sc)
rs = _settings(fb, torch, W, H)
cov = torch.from_numpy(_cov3d_np(sc["scales"], sc["rotations"])).cuda()
dp = torch.full((3, H, W), 1e-3, device="cuda")
fb_ = _fwd(ref, torch, rs, t, cov3D=cov)
gbs = [_bwd(ref, torch, rs, t, fb_, dp, 1, cov3D=cov) for _ in range(2)]
fa = _fwd(fb._C, torch, rs, t, cov3D=cov)
ga = _bwd(fb._C, torch, rs, t, fa, dp, 1, cov3D=cov)
print("lib", os.environ.get("FISHER_B200_LIB"))
for i in (0,1,2,3,4):
    a, b, b2 = ga[i].double().reshape(P,-1), gbs[0][i].double().reshape(P,-1), gbs[1][i].double().reshape(P,-1)
    d = (a-b).abs(); j = int(d.max(dim=1).values.argmax())
    print(i, "max|ours-ref|", float(d.max()), "scale", float(b.abs().max()), "ref run-to-run", float((b-b2).abs().max()), "worst row", j, "radius", int(fa[2][j]), "ours", a[j].tolist()[:3], "ref", b[j].tolist()[:3])
