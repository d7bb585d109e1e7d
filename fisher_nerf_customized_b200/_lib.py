"""ctypes binding of include/fisher_b200.h.  There is NO fallback: if the CUDA library is missing
every operator raises (the product path never routes through oracle/ or a CPU implementation)."""
import ctypes as C
import os

_PKG = os.path.dirname(os.path.abspath(__file__))
# FISHER_B200_LIB: developer override used to A/B kernel variants built by tools/variants.sh
LIB_PATH = os.environ.get("FISHER_B200_LIB") or os.path.join(_PKG, "libfisher_b200.so")

FB_OK, FB_E_INVALID, FB_E_WORKSPACE, FB_E_CUDA, FB_E_CAPACITY, FB_E_PREFILTERED = 0, -1, -2, -3, -4, -5


class fb_camera(C.Structure):
    _fields_ = [("image_width", C.c_int), ("image_height", C.c_int), ("tanfovx", C.c_float), ("tanfovy", C.c_float),
                ("scale_modifier", C.c_float), ("sh_degree", C.c_int), ("sh_coeffs", C.c_int), ("prefiltered", C.c_int),
                ("bg", C.c_float * 3), ("campos", C.c_float * 3), ("viewmatrix", C.c_float * 16),
                ("projmatrix", C.c_float * 16)]


class fb_gaussians(C.Structure):
    _fields_ = [("P", C.c_int), ("means3D", C.c_void_p), ("opacities", C.c_void_p), ("shs", C.c_void_p),
                ("colors_precomp", C.c_void_p), ("scales", C.c_void_p), ("rotations", C.c_void_p),
                ("cov3D_precomp", C.c_void_p)]


class fb_map(C.Structure):
    _fields_ = [("P", C.c_int), ("means3D", C.c_void_p), ("opacities", C.c_void_p), ("scales", C.c_void_p),
                ("rotations", C.c_void_p), ("colors", C.c_void_p)]


# name -> (restype, argtypes); the single source of truth for tests/test_abi.py as well.
vp, sz, ci, ll, cf = C.c_void_p, C.c_size_t, C.c_int, C.c_longlong, C.c_float
SIGNATURES = {
    "fb_error_string": (C.c_char_p, [ci]),
    "fb_last_cuda_error": (ci, []),
    "fb_version": (ci, []),
    "fb_launch_count": (C.c_ulonglong, []),
    "fb_profile_enable": (ci, [ci]),
    "fb_profile_classes": (ci, []),
    "fb_profile_class_name": (C.c_char_p, [ci]),
    "fb_profile_read": (ci, [vp, vp]),
    "fb_geom_bytes": (sz, [ci]),
    "fb_image_bytes": (sz, [ci, ci]),
    "fb_inst_bytes": (sz, [ci]),
    "fb_forward_scratch_bytes": (sz, [ci, ci, ci, ci]),
    "fb_backward_scratch_bytes": (sz, [ci, ci, ci, ci]),
    "fb_raster_forward_bin": (ci, [C.POINTER(fb_camera), C.POINTER(fb_gaussians), vp, sz, vp, sz, vp, sz, vp,
                                   C.POINTER(ci), vp]),
    "fb_raster_forward_render": (ci, [C.POINTER(fb_camera), C.POINTER(fb_gaussians), vp, sz, vp, sz, vp, sz, vp, sz,
                                      ci, vp, vp, vp]),
    "fb_raster_backward": (ci, [C.POINTER(fb_camera), C.POINTER(fb_gaussians), vp, vp, sz, vp, sz, vp, sz, vp, sz, ci,
                                vp, ci, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp]),
    "fb_raster_forward_render2": (ci, [C.POINTER(fb_camera), C.POINTER(fb_gaussians), vp, vp, sz, vp, sz, vp, sz, vp, sz,
                                       ci, vp, vp, vp, vp]),
    "fb_raster_backward2": (ci, [C.POINTER(fb_camera), C.POINTER(fb_gaussians), vp, vp, vp, sz, vp, sz, vp, sz, vp, sz, ci,
                                 vp, vp, ci, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp]),
    "fb_mark_visible": (ci, [ci, vp, C.POINTER(cf), C.POINTER(cf), vp, vp]),
    "fb_debug_export": (ci, [C.POINTER(fb_camera), ci, ci, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp]),
    "fb_score_workspace_bytes": (sz, [ci, ci, ci, ci, ll, ll]),
    "fb_hessian_workspace_bytes": (sz, [ci, ci, ci, ci, ll]),
    "fb_score_set_lanes": (ci, [ci]),
    "fb_score_candidates": (ci, [C.POINTER(fb_camera), C.POINTER(fb_map), vp, ci, vp, ci, cf, vp, vp, vp, sz, ci, ll, ll, vp]),
    "fb_score_candidates_sets": (ci, [C.POINTER(fb_camera), C.POINTER(fb_map), vp, ci, vp, vp, ci, cf, vp, vp, vp, sz, ci, ll, ll, vp]),
    "fb_accumulate_hessian_sets": (ci, [C.POINTER(fb_camera), C.POINTER(fb_map), vp, ci, ci, cf, vp, vp, vp, vp, sz, ci, ll, vp]),
    "fb_accumulate_hessian": (ci, [C.POINTER(fb_camera), C.POINTER(fb_map), vp, ci, ci, cf, vp, vp, vp, sz, ci, ll, vp]),
    "fb_knn_workspace_bytes": (sz, [ci]),
    "fb_dist2_knn3": (ci, [ci, vp, vp, vp, sz, vp]),
    "fb_count_visible": (ci, [ci, vp, ci, vp, vp, ci, ci, vp, vp, vp]),
    "fb_block_outer_accumulate": (ci, [ci, vp, vp, vp, vp, vp, cf, vp, vp]),
    "fb_unproject_append": (ci, [ci, ci, ci, vp, vp, vp, vp, vp, ll, vp, ci, vp, ci, vp, vp, vp, ci,
                                vp, vp]),
    "fb_morton_codes": (ci, [ci, vp, vp, vp, vp, vp]),
    "fb_occ_workspace_bytes": (sz, [ci, ci]),
    "fb_occ_update": (ci, [vp, ci, ci, vp, ci, ci, ci, vp, vp, vp, vp, cf, cf, cf, cf, vp, vp, sz, vp]),
    "fb_occ_update_static": (ci, [vp, ci, ci, vp, ci, ci, ci, vp, vp, vp, vp, cf, cf, cf, cf, vp, vp]),
    "fb_count_vs_mask": (ci, [ci, vp, vp, vp, vp, vp, ci, ci, cf, vp, vp, vp]),
    "fb_block_opt_partials": (ci, [ci]),
    "fb_block_opt": (ci, [ci, ci, vp, vp, vp, vp, cf, ci, vp, vp, vp]),
}

_lib = None


def lib():
    """Load the CUDA library once; raise loudly if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -m fisher_nerf_customized_b200.build` "
                "(there is no CPU fallback)")
        h = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(h, name)
            fn.restype = res
            fn.argtypes = args
        _lib = h
    return _lib


class FisherB200Error(RuntimeError):
    pass


def check(status, what=""):
    if status == FB_OK:
        return
    L = lib()
    msg = L.fb_error_string(status).decode()
    if status == FB_E_CUDA:
        msg += f" (cudaError {L.fb_last_cuda_error()})"
    raise FisherB200Error(f"{what}: {msg}" if what else msg)


def profile_enable(on=True):
    lib().fb_profile_enable(int(bool(on)))


def profile_read():
    """{kernel class: (total ms, launches)} since the last read (synchronises on the recorded events)."""
    L = lib()
    n = L.fb_profile_classes()
    ms, cnt = (C.c_float * n)(), (C.c_int * n)()
    check(L.fb_profile_read(ms, cnt), "fb_profile_read")
    return {L.fb_profile_class_name(i).decode(): (float(ms[i]), int(cnt[i])) for i in range(n) if cnt[i]}
