import sys; sys.path.insert(0,'.')
import numpy as np, torch
import fisher_nerf_customized_b200 as fb
from oracle import build_ref
from tests import common
from tests.golden.make_golden import _img_state
from tests.test_gpu_raster import _run_cuda, _settings, _to_t
P,W,H,seed=20000,256,256,21
sc=common.random_view_scene(P,seed,W,H); rs=_settings(fb,torch,W,H); t=_to_t(torch,sc)
got=_run_cuda(fb,torch,rs,t,0,None)
ref=build_ref.import_ref('ref_dgr'); e=torch.Tensor([])
R,color,radii,geom,binning,img,depth=ref.rasterize_gaussians(rs.bg,t['means3D'],t['rgb_colors'],t['opacities'],t['scales'],t['rotations'],1.0,e,rs.viewmatrix,rs.projmatrix,rs.tanfovx,rs.tanfovy,H,W,e,0,rs.campos,False)
T,nc,rg=_img_state(img,H*W)
print('final_T equal frac',(got['final_T']==T).mean(),'maxdiff',np.abs(got['final_T']-T).max())
print('n_contrib equal frac',(got['n_contrib']==nc).mean())
c=color.cpu().numpy(); print('color equal frac',(got['color']==c).mean())
# geometry state of reference: depths f32[P], clamped bool[3P], internal_radii i32[P], means2D float2[P], cov3D f32[6P], conic_opacity float4[P]
base=geom.data_ptr()
def al(p): return (p+127)//128*128
o=al(base)-base; depths=geom[o:o+4*P].view(torch.float32).cpu().numpy(); o=al(base+o+4*P)-base
o=al(base+o+3*P)-base if False else al(base+o+3*P)-base
ir=geom[o:o+4*P].view(torch.int32).cpu().numpy(); o=al(base+o+4*P)-base
m2=geom[o:o+8*P].view(torch.float32).cpu().numpy().reshape(P,2); o=al(base+o+8*P)-base
cov=geom[o:o+24*P].view(torch.float32).cpu().numpy(); o=al(base+o+24*P)-base
co=geom[o:o+16*P].view(torch.float32).cpu().numpy().reshape(P,4)
vis=got['radii']>0
print('radii eq',(ir==got['radii']).all(),'depth eq',(depths[vis]==got['depths'][vis]).mean(),'means2D eq',(m2[vis]==got['means2D'][vis]).mean(),'conic eq',(co[vis]==got['conic_opacity'][vis]).mean())
bad=np.where((co[vis]!=got['conic_opacity'][vis]).any(1))[0][:5]; print(co[vis][bad], got['conic_opacity'][vis][bad])
from oracle import oracle as O
from tests.test_gpu_raster import _run_oracle
orc=_run_oracle(O,rs,sc)
cov=cov.reshape(P,6)
print('ref cov3D vs oracle eq frac', (cov[vis]==orc.cov3D[vis]).mean(axis=0))
print('ref conic vs oracle eq frac', (co[vis]==orc.conic_opacity[vis]).mean(axis=0))
i=np.where((cov[vis]!=orc.cov3D[vis]).any(1))[0][:3]
print(cov[vis][i]); print(orc.cov3D[vis][i])
print('radii neq count',(ir!=got['radii']).sum(), ir[:10], got['radii'][:10])
