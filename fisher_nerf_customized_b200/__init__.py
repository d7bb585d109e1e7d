"""fisher_nerf_customized_b200 -- B200-native (sm_100a) FisherRF candidate-view scoring path.

Drop-in operator API (same names as the reference's extensions):
    GaussianRasterizer, GaussianRasterizationSettings, rasterize_gaussians   (diff_gaussian_rasterization)
    distCUDA2                                                                (simple_knn._C)
    count_visible_points                                                     (tester_navigator.py)
Additive batched API: FisherScorer, GaussianMap, ShardedScorer, count_visible_batch, setup_camera.
All compute runs in libfisher_b200.so (hand-written CUDA); there is no CPU fallback.
"""
from .rasterizer import (GaussianRasterizationSettings, GaussianRasterizer, _RasterizeGaussians,  # noqa: F401
                         rasterize_gaussians, _C, GaussianRasterizerFused)
from .knn import distCUDA2  # noqa: F401
from .visibility import count_visible_points, count_visible_batch  # noqa: F401
from .scoring import FisherScorer, GaussianMap, setup_camera  # noqa: F401
from .parallel import ShardedScorer, shard_bounds, round_robin  # noqa: F401
from .mapupdate import ResidentParams, get_pointcloud, initialize_new_params, render_rgb_depth_silhouette, get_depth_and_silhouette  # noqa: F401
from .mapnbr import OccupancyMap, count_gaussians_vs_mask  # noqa: F401
