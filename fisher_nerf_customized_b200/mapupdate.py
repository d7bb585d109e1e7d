"""Map update, densification half: new Gaussians from the masked pixels of an RGB-D frame.

Host-side mirror of (names, arguments and return values kept):
  get_pointcloud           models/SLAM/gaussian.py:75-142
  initialize_new_params    models/SLAM/gaussian.py:299-318
  add_new_gaussians' torch.cat growth  :399-401  ->  ResidentParams.append_from_frame (in place, with capacity)
One kernel (`fb_unproject_append`) replaces the ~12 full-frame temporaries + boolean selection of the reference.
There is no CPU fallback: the CUDA library must be present.
"""
import torch

from . import _lib
from .rasterizer import _f32c, _on_tensor_device, _stream


def _intr4(intrinsics, dev):
    k = torch.as_tensor(intrinsics, dtype=torch.float32).to(dev)
    return torch.stack((k[0][0], k[1][1], k[0][2], k[1][2])).contiguous()     # device, no host round trip


def _c2w16(w2c, dev):
    w2c = torch.as_tensor(w2c, dtype=torch.float32).to(dev)
    return torch.inverse(w2c).contiguous()                            # :105


def _selected_pixels(mask, height, width, downsample):
    """The reference's mask handling (:126-137): max-pool by `downsample`, flatten; indices of the set entries
    (ascending = the order boolean indexing keeps), or None for "all pixels" (no mask, or nothing left of it)."""
    if mask is None:
        return None
    m = mask.reshape(height, width)
    if downsample > 1:
        m = torch.nn.functional.max_pool2d(m.unsqueeze(0).float(), downsample).bool()
    m = m.reshape(-1)
    idx = torch.nonzero(m).reshape(-1)
    if idx.numel() == 0:
        print("[WARN] Mask become all zero after downsampling")          # :138
        return None
    return idx.contiguous()


@_on_tensor_device
def _launch(color, depth, intrinsics, w2c, downsample, pix, n, means3D, mean_stride, rgb, rgb_stride, rots=None,
            logit=None, log_scales=None, ls_cols=3, msd=None):
    L = _lib.lib()
    height, width = int(color.shape[1]), int(color.shape[2])
    p = lambda t: None if t is None else t.data_ptr()
    intr, c2w = _intr4(intrinsics, color.device), _c2w16(w2c, color.device)
    rc = L.fb_unproject_append(width, height, int(downsample), color.data_ptr(), depth.data_ptr(), intr.data_ptr(),
                               c2w.data_ptr(), p(pix), int(n), means3D.data_ptr(), mean_stride, rgb.data_ptr(), rgb_stride,
                               p(rots), p(logit), p(log_scales), ls_cols, p(msd), _stream())
    _lib.check(rc, "fb_unproject_append")


def _frame(color, depth, downsample):
    color, depth = _f32c(color), _f32c(depth)
    if not color.is_cuda or not depth.is_cuda:
        raise RuntimeError("get_pointcloud: color and depth must be CUDA tensors (there is no CPU path)")
    height, width = int(color.shape[1]), int(color.shape[2])
    if height % downsample or width % downsample:
        raise RuntimeError("get_pointcloud: image size must be a multiple of `downsample` "
                           "(the reference's grid, ones column and pooled mask disagree otherwise)")
    return color, depth, height, width


@torch.no_grad()
def get_pointcloud(color, depth, intrinsics, w2c, transform_pts=True, downsample=1, mask=None,
                   compute_mean_sq_dist=False, mean_sq_dist_method="projective"):
    """Drop-in for models/SLAM/gaussian.py:get_pointcloud: returns point_cld [N,6] (xyz + rgb) and, if asked,
    mean3_sq_dist [N]."""
    if compute_mean_sq_dist and mean_sq_dist_method != "projective":
        raise ValueError(f"Unknown mean_sq_dist_method {mean_sq_dist_method}")       # :117
    color, depth, height, width = _frame(color, depth, downsample)
    pix = _selected_pixels(mask, height, width, downsample)
    n = (height // downsample) * (width // downsample) if pix is None else int(pix.numel())
    dev = color.device
    cld = torch.empty((n, 6), dtype=torch.float32, device=dev)
    msd = torch.empty((n,), dtype=torch.float32, device=dev) if compute_mean_sq_dist else None
    eye = torch.eye(4)
    _launch(color, depth, intrinsics, w2c if transform_pts else eye, downsample, pix, n, cld, 6, cld[:, 3:], 6, msd=msd)
    return (cld, msd) if compute_mean_sq_dist else cld


class ResidentParams:
    """The raw parameter arrays of the map (the reference's `params` dict: means3D, rgb_colors, unnorm_rotations,
    logit_opacities, log_scales; gaussian.py:156-168) with CAPACITY: densification writes the new rows behind the live
    ones instead of re-allocating every array with torch.cat once per frame (:399-401, a 56-byte-per-Gaussian copy of
    the whole map).  `params()` returns views of the live rows under the reference's keys."""

    KEYS = (("means3D", 3), ("rgb_colors", 3), ("unnorm_rotations", 4), ("logit_opacities", 1), ("log_scales", 3))

    def __init__(self, capacity, isotropic=False, device="cuda"):
        self.capacity, self.isotropic, self.n = int(capacity), bool(isotropic), 0
        self.buf = {k: torch.zeros((self.capacity, (1 if (k == "log_scales" and isotropic) else c)), dtype=torch.float32,
                                   device=device) for k, c in self.KEYS}
        self.timestep = torch.zeros((self.capacity,), dtype=torch.float32, device=device)
        self._maps = []

    def attach(self, gmap):
        """Keep a GaussianMap (the activated, Morton-ordered replica the scorer reads) in step with this store: every
        append activates and appends the new rows there too (O(new rows)), so the map cannot be scored stale."""
        if gmap.P != self.n:
            raise RuntimeError(f"attach: the map holds {gmap.P} Gaussians, the store {self.n}")
        self._maps.append(gmap)
        return gmap

    def _notify(self, r0, n):
        for m in self._maps:
            m.append_rows(self.params(), r0, n)

    def params(self):
        return {k: v[:self.n] for k, v in self.buf.items()}

    def _grow(self, need):
        if need <= self.capacity:
            return
        cap = max(need, int(self.capacity * 1.5) + 1)
        for k, v in self.buf.items():
            nv = torch.zeros((cap, v.shape[1]), dtype=v.dtype, device=v.device)
            nv[:self.n] = v[:self.n]
            self.buf[k] = nv
        ts = torch.zeros((cap,), dtype=torch.float32, device=self.timestep.device)
        ts[:self.n] = self.timestep[:self.n]
        self.timestep, self.capacity = ts, cap

    @torch.no_grad()
    def append_from_frame(self, color, depth, intrinsics, w2c, mask=None, downsample=1, time_idx=0):
        """get_pointcloud(..., mask=mask, compute_mean_sq_dist=True) + initialize_new_params + the concatenation of
        add_new_gaussians (gaussian.py:353-355, 397-401, 409-410), in place.  Returns (first new row, count)."""
        color, depth, height, width = _frame(color, depth, downsample)
        pix = _selected_pixels(mask, height, width, downsample)
        n = (height // downsample) * (width // downsample) if pix is None else int(pix.numel())
        self._grow(self.n + n)
        r0, b = self.n, self.buf
        _launch(color, depth, intrinsics, w2c, downsample, pix, n, b["means3D"][r0:], 3, b["rgb_colors"][r0:], 3,
                rots=b["unnorm_rotations"][r0:], logit=b["logit_opacities"][r0:], log_scales=b["log_scales"][r0:],
                ls_cols=1 if self.isotropic else 3)
        self.timestep[r0:r0 + n] = float(time_idx)
        self.n += n
        self._notify(r0, n)
        return r0, n


@torch.no_grad()
def initialize_new_params(new_pt_cld, mean3_sq_dist, isotropic=False):
    """models/SLAM/gaussian.py:299-318 on an existing point cloud (kept for callers that build `new_pt_cld`
    themselves, e.g. with the random Gaussians of :364-393 appended); plain tensor ops, no kernel of its own."""
    n = new_pt_cld.shape[0]
    dev = new_pt_cld.device
    rots = torch.zeros((n, 4), dtype=torch.float32, device=dev)
    rots[:, 0] = 1.0
    out = {
        "means3D": new_pt_cld[:, :3],
        "rgb_colors": new_pt_cld[:, 3:6],
        "unnorm_rotations": rots,
        "logit_opacities": torch.zeros((n, 1), dtype=torch.float32, device=dev),
        "log_scales": torch.tile(torch.log(torch.sqrt(mean3_sq_dist))[..., None], (1, 3 if not isotropic else 1)),
    }
    return {k: torch.nn.Parameter(v.float().contiguous().requires_grad_(True)) for k, v in out.items()}


# ---- map update, optimisation half: the renders of get_loss in one pass -------------------------
def get_depth_and_silhouette(pts_3D, w2c):
    """slam_helpers.py:235-252: per-Gaussian colours [depth_z, 1, depth_z^2] of the depth + silhouette render."""
    pts4 = torch.cat((pts_3D, torch.ones_like(pts_3D[:, :1])), dim=-1)
    pts_in_cam = (w2c @ pts4.transpose(0, 1)).transpose(0, 1)
    depth_z = pts_in_cam[:, 2].unsqueeze(-1)
    return torch.cat((depth_z, torch.ones_like(depth_z), torch.square(depth_z)), dim=-1)


def render_rgb_depth_silhouette(params, cam, w2c, transformed_pts):
    """The two renders of get_loss / add_new_gaussians (models/SLAM/gaussian.py:203-216, 325-341) as ONE operator call:
        rendervar          = transformed_params2rendervar(params, transformed_pts)            (slam_helpers.py:178-188)
        depth_sil_rendervar = transformed_params2depthplussilhouette(params, w2c, transformed_pts)   (:268-279)
        im, radius, _ = Renderer(cam)(**rendervar);  depth_sil, _, _ = Renderer(cam)(**depth_sil_rendervar)
    Preprocess, binning and sort run once, one sweep blends the six channels, one backward sweep returns the gradients
    (differentiable through params and transformed_pts like the reference's two calls; means2D carries the colour
    render's gradient only, as variables['means2D'] does in the reference).
    Returns (im [3,H,W], radius [P], depth_sil [3,H,W], means2D [P,3] leaf whose .grad is the densification statistic)."""
    from .rasterizer import GaussianRasterizerFused
    import torch.nn.functional as F
    log_scales = params['log_scales'] if params['log_scales'].shape[-1] == 3 else torch.tile(params['log_scales'], (1, 3))
    means2D = torch.zeros_like(params['means3D'], requires_grad=True, device=params['means3D'].device) + 0
    if means2D.requires_grad:
        means2D.retain_grad()
    im, radius, _, depth_sil = GaussianRasterizerFused(cam)(
        means3D=transformed_pts, means2D=means2D, opacities=torch.sigmoid(params['logit_opacities']),
        colors_precomp=params['rgb_colors'], extra_colors=get_depth_and_silhouette(transformed_pts, w2c),
        scales=torch.exp(log_scales), rotations=F.normalize(params['unnorm_rotations']))
    return im, radius, depth_sil, means2D
