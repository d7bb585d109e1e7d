"""Shim: `from diff_gaussian_rasterization import GaussianRasterizer, GaussianRasterizationSettings`
(models/SLAM/gaussian.py:15, models/SLAM/utils/recon_helpers.py:2) resolves to the B200 implementation
when `shims/` is on sys.path ahead of the reference extension."""
from fisher_nerf_customized_b200.rasterizer import (GaussianRasterizationSettings, GaussianRasterizer,  # noqa: F401
                                                    _RasterizeGaussians, rasterize_gaussians, _C)
