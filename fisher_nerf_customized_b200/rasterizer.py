"""Drop-in operator API of `diff-gaussian-rasterization-modified` on top of libfisher_b200.so.

Mirrors (names, argument meaning, return tuples, error behaviour):
  GaussianRasterizationSettings   DGR/diff_gaussian_rasterization/__init__.py:140-151
  GaussianRasterizer              DGR/diff_gaussian_rasterization/__init__.py:153-204
  _RasterizeGaussians             DGR/diff_gaussian_rasterization/__init__.py:42-138
  rasterize_gaussians             DGR/diff_gaussian_rasterization/__init__.py:17-40
  _C.rasterize_gaussians / rasterize_gaussians_backward / mark_visible
                                  DGR/rasterize_points.h:18-66, DGR/ext.cpp:16-18
(DGR = thirdparty/diff-gaussian-rasterization-modified in the reference tree.)
"""
import ctypes as C
from typing import NamedTuple

import torch
import torch.nn as nn

from . import _lib


class GaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool


# ---- host-side camera cache: the C ABI wants the 16+16+3+3 camera floats on the host; reading
# them from CUDA tensors costs a sync, so do it once per (tensor storage, version).
_cam_cache = {}


def _host_floats(t, n):
    """Cached per tensor OBJECT (weak reference) and in-place version, never per address: a freed camera
    tensor's address is reused by the allocator."""
    import weakref
    ent = _cam_cache.get(id(t))
    if ent is not None and ent[0]() is t and ent[1] == t._version:
        return ent[2]
    v = t.detach().reshape(-1).to(dtype=torch.float32, device="cpu").tolist()
    if len(v) != n:
        raise RuntimeError(f"camera tensor must have {n} elements, got {len(v)}")
    if len(_cam_cache) > 1024:
        for k in [k for k, e in _cam_cache.items() if e[0]() is None]:
            del _cam_cache[k]
        if len(_cam_cache) > 1024:
            _cam_cache.clear()
    _cam_cache[id(t)] = (weakref.ref(t), t._version, v)
    return v


def _camera(bg, scale_modifier, viewmatrix, projmatrix, tan_fovx, tan_fovy, image_height, image_width, degree, M,
            campos, prefiltered):
    cam = _lib.fb_camera()
    cam.image_width, cam.image_height = int(image_width), int(image_height)
    cam.tanfovx, cam.tanfovy = float(tan_fovx), float(tan_fovy)
    cam.scale_modifier = float(scale_modifier)
    cam.sh_degree, cam.sh_coeffs, cam.prefiltered = int(degree), int(M), int(bool(prefiltered))
    cam.bg[:] = _host_floats(bg, 3)
    cam.campos[:] = _host_floats(campos, 3) if campos.numel() == 3 else [0.0, 0.0, 0.0]
    cam.viewmatrix[:] = _host_floats(viewmatrix, 16)
    cam.projmatrix[:] = _host_floats(projmatrix, 16)
    return cam


def _ptr(t):
    """Device pointer of an optional tensor; absent arguments arrive as empty CPU tensors
    (GaussianRasterizer.forward, __init__.py:180-190) and map to NULL."""
    if t is None or t.numel() == 0:
        return None
    return t.data_ptr()


def _f32c(t):
    if t is None or t.numel() == 0:
        return t
    if t.dtype != torch.float32:
        raise RuntimeError("fisher_b200 operators take float32 tensors")
    if not t.is_cuda:
        raise RuntimeError("fisher_b200 operators take CUDA tensors (there is no CPU path)")
    return t.contiguous()


def _stream():
    return torch.cuda.current_stream().cuda_stream


def _on_tensor_device(fn):
    """Run `fn` with the device of its first CUDA tensor argument current: the library keeps per-device state (function
    attributes, the second chunk lane) for the CURRENT device, and `_stream()` is the current device's stream."""
    import functools

    @functools.wraps(fn)
    def guarded(*args, **kw):
        for a in list(args) + list(kw.values()):
            if isinstance(a, torch.Tensor) and a.is_cuda:
                with torch.cuda.device(a.device):
                    return fn(*args, **kw)
        return fn(*args, **kw)
    return guarded


def _gauss(P, means3D, opacity, sh, colors, scales, rotations, cov3D):
    g = _lib.fb_gaussians()
    g.P = P
    g.means3D, g.opacities, g.shs, g.colors_precomp = _ptr(means3D), _ptr(opacity), _ptr(sh), _ptr(colors)
    g.scales, g.rotations, g.cov3D_precomp = _ptr(scales), _ptr(rotations), _ptr(cov3D)
    return g


class _CModule:
    """Same three entry points, argument order and return tuples as the reference's pybind `_C`."""

    @staticmethod
    @_on_tensor_device
    def rasterize_gaussians(bg, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp, viewmatrix,
                            projmatrix, tan_fovx, tan_fovy, image_height, image_width, sh, degree, campos, prefiltered):
        if means3D.ndim != 2 or means3D.size(1) != 3:
            raise RuntimeError("means3D must have dimensions (num_points, 3)")   # rasterize_points.cu:56-58
        L = _lib.lib()
        P, H, W = means3D.size(0), int(image_height), int(image_width)
        means3D, colors, opacity, scales, rotations, cov3D_precomp, sh = map(
            _f32c, (means3D, colors, opacity, scales, rotations, cov3D_precomp, sh))
        dev = means3D.device
        out_color = torch.zeros((3, H, W), dtype=torch.float32, device=dev)
        out_depth = torch.zeros((1, H, W), dtype=torch.float32, device=dev)
        radii = torch.zeros((P,), dtype=torch.int32, device=dev)
        u8 = dict(dtype=torch.uint8, device=dev)
        if P == 0:
            e = torch.empty((0,), **u8)
            return 0, out_color, radii, e, e.clone(), e.clone(), out_depth
        M = sh.size(1) if (sh is not None and sh.numel() != 0) else 0
        cam = _camera(bg, scale_modifier, viewmatrix, projmatrix, tan_fovx, tan_fovy, H, W, degree, M, campos, prefiltered)
        g = _gauss(P, means3D, opacity, sh, colors, scales, rotations, cov3D_precomp)
        geom = torch.empty((L.fb_geom_bytes(P),), **u8)
        img = torch.empty((L.fb_image_bytes(W, H),), **u8)
        n = C.c_int(0)
        st = _stream()
        _lib.check(L.fb_raster_forward_bin(C.byref(cam), C.byref(g), geom.data_ptr(), geom.numel(), img.data_ptr(),
                                           img.numel(), None, 0, radii.data_ptr(), C.byref(n), st), "rasterize_gaussians")
        R = n.value
        inst = torch.empty((L.fb_inst_bytes(R),), **u8)
        _lib.check(L.fb_raster_forward_render(C.byref(cam), C.byref(g), geom.data_ptr(), geom.numel(), img.data_ptr(),
                                              img.numel(), inst.data_ptr(), inst.numel(), None, 0, R,
                                              out_color.data_ptr(), out_depth.data_ptr(), st), "rasterize_gaussians")
        return R, out_color, radii, geom, inst, img, out_depth

    @staticmethod
    @_on_tensor_device
    def rasterize_gaussians_backward(bg, means3D, radii, colors, scales, rotations, scale_modifier, cov3D_precomp,
                                     viewmatrix, projmatrix, tan_fovx, tan_fovy, dL_dout_color, sh, degree, campos,
                                     geomBuffer, R, binningBuffer, imageBuffer, power):
        L = _lib.lib()
        P, H, W = means3D.size(0), dL_dout_color.size(1), dL_dout_color.size(2)
        M = sh.size(1) if sh.numel() != 0 else 0
        o = dict(dtype=torch.float32, device=means3D.device)
        dL_dmeans3D, dL_dmeans2D = torch.zeros((P, 3), **o), torch.zeros((P, 3), **o)
        dL_dcolors, dL_dconic = torch.zeros((P, 3), **o), torch.zeros((P, 2, 2), **o)
        dL_dopacity, dL_dcov3D = torch.zeros((P, 1), **o), torch.zeros((P, 6), **o)
        dL_dsh = torch.zeros((P, M, 3), **o)
        dL_dscales, dL_drotations = torch.zeros((P, 3), **o), torch.zeros((P, 4), **o)
        if P != 0:
            means3D, colors, scales, rotations, cov3D_precomp, sh, dL_dout_color = map(
                _f32c, (means3D, colors, scales, rotations, cov3D_precomp, sh, dL_dout_color))
            cam = _camera(bg, scale_modifier, viewmatrix, projmatrix, tan_fovx, tan_fovy, H, W, degree, M, campos, False)
            # opacities are not an input of the reference's backward (they live in its geomBuffer); ours too.
            g = _gauss(P, means3D, means3D, sh, colors, scales, rotations, cov3D_precomp)
            scratch = torch.empty((L.fb_backward_scratch_bytes(P, int(R), int(power), int(M > 0)),), dtype=torch.uint8,
                                  device=means3D.device)
            _lib.check(L.fb_raster_backward(
                C.byref(cam), C.byref(g), radii.data_ptr(), geomBuffer.data_ptr(), geomBuffer.numel(),
                imageBuffer.data_ptr(), imageBuffer.numel(), _ptr(binningBuffer), binningBuffer.numel(),
                scratch.data_ptr(), scratch.numel(), int(R), dL_dout_color.data_ptr(), int(power),
                dL_dmeans2D.data_ptr(), dL_dconic.data_ptr(), dL_dopacity.data_ptr(), dL_dcolors.data_ptr(),
                dL_dmeans3D.data_ptr(), dL_dcov3D.data_ptr(), _ptr(dL_dsh), dL_dscales.data_ptr(),
                dL_drotations.data_ptr(), _stream()), "rasterize_gaussians_backward")
        return dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations

    @staticmethod
    @_on_tensor_device
    def rasterize_gaussians2(bg, means3D, colors, extra_colors, opacity, scales, rotations, scale_modifier, viewmatrix,
                             projmatrix, tan_fovx, tan_fovy, image_height, image_width, campos, prefiltered):
        """Colour render and a second render of `extra_colors` [P,3] over the same geometry in one pass:
        returns (num_rendered, color, extra, radii, geomBuffer, binningBuffer, imgBuffer, depth)."""
        if means3D.ndim != 2 or means3D.size(1) != 3:
            raise RuntimeError("means3D must have dimensions (num_points, 3)")
        L = _lib.lib()
        P, H, W = means3D.size(0), int(image_height), int(image_width)
        means3D, colors, extra_colors, opacity, scales, rotations = map(
            _f32c, (means3D, colors, extra_colors, opacity, scales, rotations))
        dev = means3D.device
        out_color = torch.zeros((3, H, W), dtype=torch.float32, device=dev)
        out_extra = torch.zeros((3, H, W), dtype=torch.float32, device=dev)
        out_depth = torch.zeros((1, H, W), dtype=torch.float32, device=dev)
        radii = torch.zeros((P,), dtype=torch.int32, device=dev)
        u8 = dict(dtype=torch.uint8, device=dev)
        if P == 0:
            e = torch.empty((0,), **u8)
            return 0, out_color, out_extra, radii, e, e.clone(), e.clone(), out_depth
        cam = _camera(bg, scale_modifier, viewmatrix, projmatrix, tan_fovx, tan_fovy, H, W, 0, 0, campos, prefiltered)
        g = _gauss(P, means3D, opacity, None, colors, scales, rotations, None)
        geom = torch.empty((L.fb_geom_bytes(P),), **u8)
        img = torch.empty((L.fb_image_bytes(W, H),), **u8)
        n = C.c_int(0)
        st = _stream()
        _lib.check(L.fb_raster_forward_bin(C.byref(cam), C.byref(g), geom.data_ptr(), geom.numel(), img.data_ptr(),
                                           img.numel(), None, 0, radii.data_ptr(), C.byref(n), st), "rasterize_gaussians2")
        R = n.value
        inst = torch.empty((L.fb_inst_bytes(R),), **u8)
        _lib.check(L.fb_raster_forward_render2(C.byref(cam), C.byref(g), extra_colors.data_ptr(), geom.data_ptr(),
                                               geom.numel(), img.data_ptr(), img.numel(), inst.data_ptr(), inst.numel(),
                                               None, 0, R, out_color.data_ptr(), out_extra.data_ptr(),
                                               out_depth.data_ptr(), st), "rasterize_gaussians2")
        return R, out_color, out_extra, radii, geom, inst, img, out_depth

    @staticmethod
    @_on_tensor_device
    def rasterize_gaussians_backward2(bg, means3D, radii, colors, extra_colors, scales, rotations, scale_modifier,
                                      viewmatrix, projmatrix, tan_fovx, tan_fovy, dL_dout_color, dL_dout_extra, campos,
                                      geomBuffer, R, binningBuffer, imageBuffer):
        """Backward of rasterize_gaussians2 (power 1): (dL_dmeans2D of the colour render alone, dL_dcolors, dL_dextra,
        dL_dopacity, dL_dmeans3D, dL_dscales, dL_drotations), the shared inputs' gradients summed over both renders."""
        L = _lib.lib()
        P, H, W = means3D.size(0), dL_dout_color.size(1), dL_dout_color.size(2)
        o = dict(dtype=torch.float32, device=means3D.device)
        dL_dmeans3D, dL_dmeans2D = torch.zeros((P, 3), **o), torch.zeros((P, 3), **o)
        dL_dcolors, dL_dextra = torch.zeros((P, 3), **o), torch.zeros((P, 3), **o)
        dL_dopacity = torch.zeros((P, 1), **o)
        dL_dscales, dL_drotations = torch.zeros((P, 3), **o), torch.zeros((P, 4), **o)
        if P != 0 and int(R) != 0:
            means3D, colors, extra_colors, scales, rotations, dL_dout_color, dL_dout_extra = map(
                _f32c, (means3D, colors, extra_colors, scales, rotations, dL_dout_color, dL_dout_extra))
            cam = _camera(bg, scale_modifier, viewmatrix, projmatrix, tan_fovx, tan_fovy, H, W, 0, 0, campos, False)
            g = _gauss(P, means3D, means3D, None, colors, scales, rotations, None)
            scratch = torch.empty((L.fb_backward_scratch_bytes(P, int(R), 1, 0),), dtype=torch.uint8, device=means3D.device)
            _lib.check(L.fb_raster_backward2(
                C.byref(cam), C.byref(g), extra_colors.data_ptr(), radii.data_ptr(), geomBuffer.data_ptr(),
                geomBuffer.numel(), imageBuffer.data_ptr(), imageBuffer.numel(), _ptr(binningBuffer), binningBuffer.numel(),
                scratch.data_ptr(), scratch.numel(), int(R), dL_dout_color.data_ptr(), dL_dout_extra.data_ptr(), 1,
                dL_dmeans2D.data_ptr(), None, dL_dopacity.data_ptr(), dL_dcolors.data_ptr(), dL_dextra.data_ptr(),
                dL_dmeans3D.data_ptr(), None, None, dL_dscales.data_ptr(), dL_drotations.data_ptr(), _stream()),
                "rasterize_gaussians_backward2")
        return dL_dmeans2D, dL_dcolors, dL_dextra, dL_dopacity, dL_dmeans3D, dL_dscales, dL_drotations

    @staticmethod
    @_on_tensor_device
    def mark_visible(means3D, viewmatrix, projmatrix):
        L = _lib.lib()
        P = means3D.size(0)
        present = torch.zeros((P,), dtype=torch.bool, device=means3D.device)
        if P != 0:
            vm = (C.c_float * 16)(*_host_floats(viewmatrix, 16))
            pm = (C.c_float * 16)(*_host_floats(projmatrix, 16))
            _lib.check(L.fb_mark_visible(P, _f32c(means3D).data_ptr(), vm, pm, present.data_ptr(), _stream()),
                       "mark_visible")
        return present


_C = _CModule()


class _RasterizeGaussians(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                raster_settings, backward_power):
        rs = raster_settings
        num_rendered, color, radii, geomBuffer, binningBuffer, imgBuffer, depth = _C.rasterize_gaussians(
            rs.bg, means3D, colors_precomp, opacities, scales, rotations, rs.scale_modifier, cov3Ds_precomp,
            rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, rs.image_height, rs.image_width, sh, rs.sh_degree,
            rs.campos, rs.prefiltered)
        ctx.raster_settings = rs
        ctx.num_rendered = num_rendered
        ctx.backward_power = backward_power
        ctx.save_for_backward(colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, geomBuffer,
                              binningBuffer, imgBuffer)
        ctx.mark_non_differentiable(radii)
        return color, radii, depth

    @staticmethod
    def backward(ctx, grad_out_color, _grad_radii, _grad_depth):
        # gradients w.r.t. radii and depth are ignored, as in the reference (__init__.py:92)
        rs = ctx.raster_settings
        colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, geomBuffer, binningBuffer, imgBuffer = \
            ctx.saved_tensors
        (grad_means2D, grad_colors_precomp, grad_opacities, grad_means3D, grad_cov3Ds_precomp, grad_sh, grad_scales,
         grad_rotations) = _C.rasterize_gaussians_backward(
            rs.bg, means3D, radii, colors_precomp, scales, rotations, rs.scale_modifier, cov3Ds_precomp, rs.viewmatrix,
            rs.projmatrix, rs.tanfovx, rs.tanfovy, grad_out_color, sh, rs.sh_degree, rs.campos, geomBuffer,
            ctx.num_rendered, binningBuffer, imgBuffer, ctx.backward_power)
        return (grad_means3D, grad_means2D, grad_sh, grad_colors_precomp, grad_opacities, grad_scales, grad_rotations,
                grad_cov3Ds_precomp, None, None)


def rasterize_gaussians(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                        raster_settings, backward_power):
    return _RasterizeGaussians.apply(means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                                     cov3Ds_precomp, raster_settings, backward_power)


class _RasterizeGaussians2(torch.autograd.Function):
    """Autograd bridge of the fused colour + extra render (inputs in the order of the reference's bridge, with the
    second colour triple after the first)."""

    @staticmethod
    def forward(ctx, means3D, means2D, colors_precomp, extra_colors, opacities, scales, rotations, raster_settings):
        rs = raster_settings
        num_rendered, color, extra, radii, geomBuffer, binningBuffer, imgBuffer, depth = _C.rasterize_gaussians2(
            rs.bg, means3D, colors_precomp, extra_colors, opacities, scales, rotations, rs.scale_modifier, rs.viewmatrix,
            rs.projmatrix, rs.tanfovx, rs.tanfovy, rs.image_height, rs.image_width, rs.campos, rs.prefiltered)
        ctx.raster_settings = rs
        ctx.num_rendered = num_rendered
        ctx.save_for_backward(colors_precomp, extra_colors, means3D, scales, rotations, radii, geomBuffer, binningBuffer,
                              imgBuffer)
        ctx.mark_non_differentiable(radii, depth)
        return color, radii, depth, extra

    @staticmethod
    def backward(ctx, grad_color, _grad_radii, _grad_depth, grad_extra):
        rs = ctx.raster_settings
        colors_precomp, extra_colors, means3D, scales, rotations, radii, geomBuffer, binningBuffer, imgBuffer = ctx.saved_tensors
        H, W = rs.image_height, rs.image_width
        if grad_color is None:
            grad_color = torch.zeros((3, H, W), dtype=torch.float32, device=means3D.device)
        if grad_extra is None:
            grad_extra = torch.zeros((3, H, W), dtype=torch.float32, device=means3D.device)
        g2d, gcol, gext, gop, g3d, gsc, grot = _C.rasterize_gaussians_backward2(
            rs.bg, means3D, radii, colors_precomp, extra_colors, scales, rotations, rs.scale_modifier, rs.viewmatrix,
            rs.projmatrix, rs.tanfovx, rs.tanfovy, grad_color, grad_extra, rs.campos, geomBuffer, ctx.num_rendered,
            binningBuffer, imgBuffer)
        return g3d, g2d, gcol, gext, gop, gsc, grot, None


class GaussianRasterizerFused(nn.Module):
    """Two renders of one geometry in one pass: `forward(...)` returns (color [3,H,W], radii, depth, extra [3,H,W]), the
    results of `GaussianRasterizer(rs)(colors_precomp=colors)` and `GaussianRasterizer(rs)(colors_precomp=extra)[0]` --
    what get_loss / add_new_gaussians do with two operator calls (models/SLAM/gaussian.py:207,211,325-341).
    backward_power is 1 (the two renders' gradients add); means2D receives the colour render's gradient only."""

    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def forward(self, means3D, means2D, opacities, colors_precomp, extra_colors, scales, rotations):
        return _RasterizeGaussians2.apply(means3D, means2D, colors_precomp, extra_colors, opacities, scales, rotations,
                                          self.raster_settings)


class GaussianRasterizer(nn.Module):
    def __init__(self, raster_settings, backward_power: int = 1):
        super().__init__()
        self.raster_settings = raster_settings
        self.backward_power = backward_power

    def markVisible(self, positions):
        with torch.no_grad():
            rs = self.raster_settings
            return _C.mark_visible(positions, rs.viewmatrix, rs.projmatrix)

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                cov3D_precomp=None):
        if (shs is None) == (colors_precomp is None):
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')
        has_sr = scales is not None or rotations is not None
        if ((scales is None or rotations is None) and cov3D_precomp is None) or (has_sr and cov3D_precomp is not None):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')
        empty = torch.Tensor([])
        shs = empty if shs is None else shs
        colors_precomp = empty if colors_precomp is None else colors_precomp
        scales = empty if scales is None else scales
        rotations = empty if rotations is None else rotations
        cov3D_precomp = empty if cov3D_precomp is None else cov3D_precomp
        return rasterize_gaussians(means3D, means2D, shs, colors_precomp, opacities, scales, rotations, cov3D_precomp,
                                   self.raster_settings, self.backward_power)


def debug_export(raster_settings, P, num_rendered, geomBuffer, binningBuffer, imgBuffer):
    """Parity-test introspection (fb_debug_export): the private per-call state translated into the
    reference's vocabulary (GeometryState / BinningState / ImageState, rasterizer_impl.h:30-70) as numpy."""
    rs = raster_settings
    L = _lib.lib()
    H, W = int(rs.image_height), int(rs.image_width)
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    dev = geomBuffer.device
    cam = _camera(rs.bg, rs.scale_modifier, rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, H, W, rs.sh_degree, 0,
                  rs.campos, rs.prefiltered)
    R = int(num_rendered)
    out = dict(tiles_touched=torch.zeros(P, dtype=torch.int32, device=dev), depths=torch.zeros(P, device=dev),
               means2D=torch.zeros((P, 2), device=dev), conic_opacity=torch.zeros((P, 4), device=dev),
               ranges=torch.zeros((tiles, 2), dtype=torch.int32, device=dev), final_T=torch.zeros(H * W, device=dev),
               n_contrib=torch.zeros(H * W, dtype=torch.int32, device=dev),
               point_list=torch.zeros(max(R, 1), dtype=torch.int32, device=dev),
               point_keys=torch.zeros(max(R, 1), dtype=torch.int64, device=dev))
    _lib.check(L.fb_debug_export(C.byref(cam), P, R, geomBuffer.data_ptr(), imgBuffer.data_ptr(),
                                 _ptr(binningBuffer) if R > 0 else None, out["tiles_touched"].data_ptr(),
                                 out["depths"].data_ptr(), out["means2D"].data_ptr(), out["conic_opacity"].data_ptr(),
                                 out["ranges"].data_ptr(), out["final_T"].data_ptr(), out["n_contrib"].data_ptr(),
                                 out["point_list"].data_ptr(), out["point_keys"].data_ptr(), _stream()), "debug_export")
    res = {k: v.cpu().numpy() for k, v in out.items()}
    res["tiles_touched"] = res["tiles_touched"].view("uint32")
    res["ranges"] = res["ranges"].view("uint32")
    res["n_contrib"] = res["n_contrib"].view("uint32")
    res["point_list"] = res["point_list"].view("uint32")[:R]
    res["point_keys"] = res["point_keys"].view("uint64")[:R]
    return res
