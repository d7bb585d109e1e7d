"""Occupancy grid and masked visibility counter, host-side mirrors of

  AstarPlanner.init / update_occ_map      planning/astar.py:64-106, 202-303
  count_gaussians_vs_mask                 models/SLAM/gaussian_object.py:359-428

The reference's update_occ_map sorts ~11 * H * W rows twice with torch.unique, copies the occupied cells to the
host and draws one cv2.line per cell in a Python loop.  Here the whole update is three kernels behind
`fb_occ_update` and the pose never leaves the device.  There is no CPU fallback.
"""
import numpy as np
import torch

from . import _lib
from .rasterizer import _f32c, _on_tensor_device, _stream


class OccupancyMap:
    """The occ_map state of AstarPlanner (channels: unknown, occupied, free) and its per-step update."""

    def __init__(self, cell_size, height_lower, height_upper, pcd_far_distance, device="cuda"):
        self.cell_size = float(cell_size)
        self.height_lower, self.height_upper = float(height_lower), float(height_upper)
        self.pcd_far_distance = float(pcd_far_distance)
        self.device = torch.device(device)
        self.frame_idx = 0

    def init(self, pose, intrinsic, scene_bounds=None):
        """AstarPlanner.init (astar.py:64-106): fixed 768 x 768 grid centred on the first pose, or sized by the bounds."""
        pose = np.asarray(pose.detach().cpu().numpy() if torch.is_tensor(pose) else pose)
        self.grid_dim = np.array([768, 768])
        self.intrinsics = intrinsic
        self.cam_height = pose[1, 3]
        if scene_bounds is not None:
            lo, hi = scene_bounds
            centre = (hi[[0, 2]] + lo[[0, 2]]) / 2
            self.grid_dim = np.array([int((hi[0] - lo[0]) / self.cell_size + 1), int((hi[2] - lo[2]) / self.cell_size + 1)])
        else:
            centre = pose[[0, 2], 3]
        gx, gz = int(self.grid_dim[0]), int(self.grid_dim[1])
        self.occ_map = torch.zeros((3, gz, gx), device=self.device)
        self.occ_map[0] = 1.
        cx = int((pose[0, 3] - centre[0]) / self.cell_size + gx // 2)
        cz = int((pose[2, 3] - centre[1]) / self.cell_size + gz // 2)
        self.occ_map[2, cz - 1:cz + 2, cx - 1:cx + 2] = 2.
        self.map_center = torch.as_tensor(np.asarray(centre, dtype=np.float32), device=self.device)
        self._cam_cell = torch.tensor([cx, cz], dtype=torch.int32, device=self.device)
        k = torch.as_tensor(np.asarray(intrinsic.detach().cpu().numpy() if torch.is_tensor(intrinsic) else intrinsic),
                            dtype=torch.float32)
        self._intr = torch.stack((k[0, 0], k[1, 1], k[0, 2], k[1, 2])).to(self.device).contiguous()
        planes = torch.linspace(1e-3, 0.95, 11, device=self.device).clamp_(min=0.)          # astar.py:237-238
        planes[-1] = 1.                                                                      # :239 (broadcast plane)
        self._planes = planes.contiguous()
        L = _lib.lib()
        self._ws = torch.empty((L.fb_occ_workspace_bytes(gx, gz),), dtype=torch.uint8, device=self.device)
        self.frame_idx = 0

    @property
    def cam_pos(self):
        """[z, x] cell of the camera (the reference keeps it as a numpy array, astar.py:213); reading it syncs."""
        x, z = self._cam_cell.tolist()
        return np.array([z, x])

    @torch.no_grad()
    @_on_tensor_device
    def update_occ_map(self, depth, c2w, t, downsample=1):
        """Update the occupancy map from a depth observation [1,H,W] at pose c2w (astar.py:202-303)."""
        self.frame_idx = t
        if isinstance(depth, np.ndarray):
            depth = torch.from_numpy(depth)
        depth = _f32c(depth.to(self.device).float())
        c2w = _f32c(torch.as_tensor(c2w).to(self.device).float())
        H, W = int(depth.shape[1]), int(depth.shape[2])
        gx, gz = int(self.grid_dim[0]), int(self.grid_dim[1])
        L = _lib.lib()
        rc = L.fb_occ_update(self.occ_map.data_ptr(), gx, gz, depth.data_ptr(), W, H, int(downsample), c2w.data_ptr(),
                             self._intr.data_ptr(), self._planes.data_ptr(), self.map_center.data_ptr(), self.cell_size,
                             self.height_lower, self.height_upper, self.pcd_far_distance, self._cam_cell.data_ptr(),
                             self._ws.data_ptr(), self._ws.numel(), _stream())
        _lib.check(rc, "fb_occ_update")


@torch.no_grad()
@_on_tensor_device
def count_gaussians_vs_mask(params, w2c, intrinsics, obj_mask_2d, opacity_thresh=0.01):
    """Drop-in for models/SLAM/gaussian_object.py:359-428: same arguments, same dict ('in', 'out', 'active' as Python
    ints, 'idx_in', 'idx_out' as index tensors in the reference's order)."""
    N = params['means3D'].shape[0]
    if obj_mask_2d is None:
        return {'in': 0, 'out': N, 'active': 0, 'idx_in': torch.empty(0, dtype=torch.long), 'idx_out': torch.arange(N)}
    means = _f32c(params['means3D'].detach())
    dev = means.device
    H, W = int(obj_mask_2d.shape[-2]), int(obj_mask_2d.shape[-1])
    mask = obj_mask_2d.to(dev).bool().reshape(H, W).to(torch.uint8).contiguous()
    logit = _f32c(params['logit_opacities'].detach().reshape(-1))
    k = intrinsics
    intr = torch.tensor([float(k[0][0]), float(k[1][1]), float(k[0][2]), float(k[1][2])], dtype=torch.float32, device=dev)
    w = _f32c(torch.as_tensor(w2c).to(dev).float())
    code = torch.empty((N,), dtype=torch.uint8, device=dev)
    counts = torch.zeros((3,), dtype=torch.int32, device=dev)
    L = _lib.lib()
    rc = L.fb_count_vs_mask(N, means.data_ptr(), logit.data_ptr(), w.data_ptr(), intr.data_ptr(), mask.data_ptr(), W, H,
                            float(opacity_thresh), code.data_ptr(), counts.data_ptr(), _stream())
    _lib.check(rc, "fb_count_vs_mask")
    n_in, n_out, n_act = counts.tolist()
    if n_act == 0:
        e = torch.empty(0, dtype=torch.long, device=dev)
        return {'in': 0, 'out': 0, 'active': 0, 'idx_in': e, 'idx_out': e.clone()}
    idx_in = torch.nonzero(code == 1).squeeze(1)
    idx_out = torch.cat([torch.nonzero(code == 2).squeeze(1), torch.nonzero(code == 3).squeeze(1)], dim=0)
    return {'in': n_in, 'out': n_out, 'active': n_act, 'idx_in': idx_in, 'idx_out': idx_out}
