"""`count_visible_points` on libfisher_b200.so.

Mirrors tester_navigator.py:147-180: transform the N points by inv(c2w), keep z > 0, pinhole-project,
keep 0 <= u < W and 0 <= v < H (W = H = img_size), all in float64 as numpy does; returns the
visible points, in input order.  `count_visible_batch` is the additive batched form (one pass over
the points for all K poses, integer counts only) used by the active-mapping loop.
"""
import numpy as np
import torch

from . import _lib
from .rasterizer import _on_tensor_device, _stream


def _intr(intrinsics):
    k = np.asarray(intrinsics.detach().cpu().numpy() if torch.is_tensor(intrinsics) else intrinsics)
    # the reference multiplies float64 arrays by these scalars: float32 intrinsics are widened exactly
    return [float(k[0, 0]), float(k[1, 1]), float(k[0, 2]), float(k[1, 2])]


def _w2c64(c2w):
    c2w = np.asarray(c2w.detach().cpu().numpy() if torch.is_tensor(c2w) else c2w)
    c2w = c2w.reshape(-1, 4, 4)
    # np.linalg.inv on the host, exactly the reference's call (float32 input -> float32 inverse, widened after)
    return np.stack([np.linalg.inv(m) for m in c2w]).astype(np.float64)


@_on_tensor_device
def count_visible_batch(points, c2w, intrinsics, img_size, return_mask=False):
    """points: f32 [N,3] CUDA tensor; c2w: [K,4,4] (numpy or tensor).  Returns int32 counts [K] on the
    device (and, optionally, the bool mask [K,N])."""
    if not (torch.is_tensor(points) and points.is_cuda):
        raise RuntimeError("count_visible_batch takes a CUDA point tensor (there is no CPU path)")
    L = _lib.lib()
    dev = points.device
    pts = points.contiguous().float()
    N = pts.size(0)
    w2c = torch.from_numpy(np.ascontiguousarray(_w2c64(c2w))).to(dev)
    K = w2c.size(0)
    intr = torch.tensor(_intr(intrinsics), dtype=torch.float64, device=dev)
    H = W = int(img_size)
    counts = torch.zeros((K,), dtype=torch.int32, device=dev)
    mask = torch.zeros((K, N), dtype=torch.uint8, device=dev) if return_mask else None
    for k0 in range(0, K, 256):
        k1 = min(K, k0 + 256)
        _lib.check(L.fb_count_visible(N, pts.data_ptr() if N else None, k1 - k0, w2c[k0:k1].data_ptr(), intr.data_ptr(),
                                      W, H, counts[k0:k1].data_ptr(),
                                      mask[k0:k1].data_ptr() if mask is not None else None, _stream()),
                   "count_visible_points")
    return (counts, mask.bool()) if return_mask else counts


def count_visible_points(global_pts, c2w, intrinsics, img_size):
    """Drop-in: same arguments and return value (the visible subset of global_pts) as the reference.
    numpy input is uploaded, the subset comes back as numpy; CUDA tensor input stays on the device."""
    is_np = not torch.is_tensor(global_pts)
    pts = torch.as_tensor(global_pts).cuda() if is_np else global_pts
    _, mask = count_visible_batch(pts, c2w, intrinsics, img_size, return_mask=True)
    vis = pts[mask[0]]
    if is_np:
        return vis.cpu().numpy().astype(np.asarray(global_pts).dtype, copy=False)
    return vis
