"""Batched FisherRF candidate-view scoring against a resident Gaussian map.

Host-side mirror of the reference callers (names and arithmetic kept):
  setup_camera                    models/SLAM/utils/recon_helpers.py:4-32
  GaussianSLAM.compute_Hessian    models/SLAM/gaussian.py:1503-1570   (C_H = 4)
  GaussianObjectSLAM.compute_Hessian  models/SLAM/gaussian_object.py:1940-2045  (C_H = 11)
  GaussianSLAM.compute_H_train    models/SLAM/gaussian.py:1338-1348
  GaussianSLAM.pose_eval          models/SLAM/gaussian.py:1354-1375
The K-iteration Python loop with one host sync per pose becomes one launch sequence per chunk of
poses (`fb_score_candidates`), with no host sync until the scores are read.
"""
import ctypes as C

import torch

from . import _lib
from .rasterizer import GaussianRasterizationSettings, _camera, _f32c, _on_tensor_device, _stream


def setup_camera(w, h, k, w2c, near=0.01, far=100, device="cuda"):
    """Same camera as the reference's setup_camera (OpenGL-style projection, matrices stored transposed)."""
    fx, fy, cx, cy = float(k[0][0]), float(k[1][1]), float(k[0][2]), float(k[1][2])
    w2c = torch.as_tensor(w2c, dtype=torch.float32).to(device)
    cam_center = torch.inverse(w2c)[:3, 3]
    w2c_t = w2c.unsqueeze(0).transpose(1, 2)
    proj = torch.tensor([[2 * fx / w, 0.0, -(w - 2 * cx) / w, 0.0],
                         [0.0, 2 * fy / h, -(h - 2 * cy) / h, 0.0],
                         [0.0, 0.0, far / (far - near), -(far * near) / (far - near)],
                         [0.0, 0.0, 1.0, 0.0]], dtype=torch.float32, device=device).unsqueeze(0).transpose(1, 2)
    full_proj = w2c_t.bmm(proj)
    return GaussianRasterizationSettings(
        image_height=h, image_width=w, tanfovx=w / (2 * fx), tanfovy=h / (2 * fy),
        bg=torch.tensor([0, 0, 0], dtype=torch.float32, device=device), scale_modifier=1.0,
        viewmatrix=w2c_t, projmatrix=full_proj, sh_degree=0, campos=cam_center, prefiltered=False)


def read_npz_params(path, device="cuda"):
    """The raw parameter dict of a reference checkpoint, `params.npz` / `params{t}.npz` as written by save_params /
    save_params_ckpt (models/SLAM/utils/common_utils.py:30-59; loaded by tester_gaussians_navigation.py:2745-2762).
    Camera tracks and the extras stored beside the parameters are ignored."""
    import numpy as np
    keys = ("means3D", "rgb_colors", "unnorm_rotations", "logit_opacities", "log_scales")
    with np.load(path, allow_pickle=True) as z:
        missing = [k for k in keys if k not in z.files]
        if missing:
            raise KeyError(f"{path}: not a reference parameter checkpoint, missing {missing}")
        return {k: torch.as_tensor(np.asarray(z[k]), dtype=torch.float32, device=device) for k in keys}


def morton_order(means3D, bits=10):
    """Permutation that sorts points by the 30-bit Morton code of their position in the map's bounding
    box (one kernel for the codes + one device radix sort, no host sync): Gaussians that are close in space become close
    in memory, so the 32 Gaussians of a warp land in a handful of tiles (warp-aggregated counters) and a tile's list
    gathers neighbouring records."""
    assert bits == 10
    means3D = _f32c(means3D)
    P = means3D.shape[0]
    lo = means3D.amin(dim=0).contiguous()
    ext = (means3D.amax(dim=0) - lo).contiguous()
    codes = torch.empty((P,), dtype=torch.int32, device=means3D.device)
    with torch.cuda.device(means3D.device):
        _lib.check(_lib.lib().fb_morton_codes(P, means3D.data_ptr(), lo.data_ptr(), ext.data_ptr(), codes.data_ptr(),
                                              _stream()), "fb_morton_codes")
    return torch.argsort(codes, stable=True)


class GaussianMap:
    """The resident map in the activated form compute_Hessian feeds the rasteriser
    (gaussian.py:1523-1543): world-frame means, normalised rotations, sigmoid opacities, exp scales, rgb.

    HBM layout: with `spatial_order=True` (default) the replica is stored in Morton order of the means
    (`perm[i]` = caller's index of the Gaussian stored at row i).  Everything that crosses the API in per-Gaussian
    form (H_train, H_train_inv, compute_Hessian) stays in the CALLER's order; FisherScorer permutes at the
    boundary (one 4*C_H*P-byte gather per call).  Scores are sums over Gaussians and do not depend on the order.

    The arrays have CAPACITY: `append_rows` activates only the rows a densification step added and writes them behind
    the live rows (unsorted tail: the order is a layout, not a result); when the tail outgrows `resort_fraction` of the
    map the whole replica is re-sorted once.  `ResidentParams.attach(map)` keeps an attached map in step with the raw
    parameter store, so a map can never be scored stale."""

    NAMES = ("means3D", "colors", "rotations", "opacities", "scales")

    def __init__(self, means3D, rgb_colors, rotations, opacities, scales, spatial_order=True, capacity=None,
                 resort_fraction=0.05):
        scales = scales if scales.shape[-1] == 3 else scales.repeat(1, 3)   # isotropic: gaussian.py:1532-1533
        self.spatial_order = bool(spatial_order)
        self.resort_fraction = float(resort_fraction)
        self.perm = None
        self._inv_perm = None
        self.tail = 0                                                        # rows appended since the last sort
        with torch.no_grad():
            if self.spatial_order and means3D.shape[0] > 1:
                self.perm = morton_order(means3D.detach().float())
                pick = lambda t: t.detach().index_select(0, self.perm)
                means3D, rgb_colors, rotations = pick(means3D), pick(rgb_colors), pick(rotations)
                opacities, scales = pick(opacities.reshape(-1)), pick(scales)
            rows = dict(means3D=_f32c(means3D.detach()), colors=_f32c(rgb_colors.detach()), rotations=_f32c(rotations.detach()),
                        opacities=_f32c(opacities.detach().reshape(-1)), scales=_f32c(scales.detach()))
            self.P = rows["means3D"].shape[0]
            self.capacity = max(int(capacity or 0), self.P)
            self._buf = {}
            for k, v in rows.items():
                if self.capacity > self.P:
                    b = torch.empty((self.capacity,) + tuple(v.shape[1:]), dtype=torch.float32, device=v.device)
                    b[:self.P] = v
                else:
                    b = v
                self._buf[k] = b

    # live rows under the names the kernels' fb_map uses
    means3D = property(lambda self: self._buf["means3D"][:self.P])
    colors = property(lambda self: self._buf["colors"][:self.P])
    rotations = property(lambda self: self._buf["rotations"][:self.P])
    opacities = property(lambda self: self._buf["opacities"][:self.P])
    scales = property(lambda self: self._buf["scales"][:self.P])

    def inverse_perm(self):
        """inverse_perm()[c] = resident row of the caller's Gaussian c (built once, on first use)."""
        if self._inv_perm is None:
            inv = torch.empty_like(self.perm)
            inv[self.perm] = torch.arange(self.P, device=self.perm.device)
            self._inv_perm = inv
        return self._inv_perm

    def to_map_order(self, per_gaussian):
        """[P, ...] tensor in the caller's order -> the resident (Morton) order."""
        return per_gaussian if self.perm is None else per_gaussian.index_select(0, self.perm)

    def to_caller_order(self, per_gaussian):
        if self.perm is None:
            return per_gaussian
        out = torch.empty_like(per_gaussian)
        out.index_copy_(0, self.perm, per_gaussian)
        return out

    @torch.no_grad()
    def append_rows(self, params, r0, n):
        """Rows [r0, r0 + n) of the reference's raw parameter dict were appended behind the caller's existing P rows
        (r0 == P): activate them (gaussian.py:1529-1533) and write them behind the live rows.  O(n), not O(P)."""
        r0, n = int(r0), int(n)
        if r0 != self.P:
            raise RuntimeError(f"append_rows: the map holds {self.P} Gaussians, the new rows start at {r0}")
        if n == 0:
            return
        sl = slice(r0, r0 + n)
        ls = params["log_scales"][sl]
        new = dict(means3D=params["means3D"][sl], colors=params["rgb_colors"][sl],
                   rotations=torch.nn.functional.normalize(params["unnorm_rotations"][sl]),
                   opacities=torch.sigmoid(params["logit_opacities"][sl]).reshape(-1),
                   scales=torch.exp(ls if ls.shape[-1] == 3 else ls.repeat(1, 3)))
        need = self.P + n
        if need > self.capacity:
            cap = max(need, int(self.capacity * 1.5) + 1)
            for k, b in self._buf.items():
                nb = torch.empty((cap,) + tuple(b.shape[1:]), dtype=torch.float32, device=b.device)
                nb[:self.P] = b[:self.P]
                self._buf[k] = nb
            self.capacity = cap
        for k, v in new.items():
            self._buf[k][self.P:need] = v.detach().float()
        if self.perm is not None:
            self.perm = torch.cat([self.perm, torch.arange(r0, r0 + n, device=self.perm.device, dtype=self.perm.dtype)])
        elif self.spatial_order and need > 1:
            self.perm = torch.arange(need, device=self._buf["means3D"].device)
        self._inv_perm = None
        self.P = need
        self.tail += n
        if self.spatial_order and self.tail > max(4096, int(self.resort_fraction * self.P)):
            self.resort()

    @torch.no_grad()
    def resort(self):
        """Re-establish the Morton order of the whole replica (after enough appended rows): one code kernel, one sort,
        five gathers; the caller-order bookkeeping (`perm`) follows."""
        if not self.spatial_order or self.P < 2:
            return
        order = morton_order(self.means3D)
        for k, b in self._buf.items():
            nb = torch.empty_like(b)
            nb[:self.P] = b[:self.P].index_select(0, order)
            self._buf[k] = nb
        self.perm = order if self.perm is None else self.perm.index_select(0, order)
        self._inv_perm = None
        self.tail = 0

    @classmethod
    def from_params(cls, params, spatial_order=True, capacity=None):
        """From the reference's raw parameter dict (gaussian.py:156-168)."""
        with torch.no_grad():
            return cls(params["means3D"], params["rgb_colors"],
                       torch.nn.functional.normalize(params["unnorm_rotations"]),
                       torch.sigmoid(params["logit_opacities"]), torch.exp(params["log_scales"]),
                       spatial_order=spatial_order, capacity=capacity)

    @classmethod
    def from_npz(cls, path, device="cuda", spatial_order=True):
        """From a reference checkpoint (see read_npz_params)."""
        return cls.from_params(read_npz_params(path, device), spatial_order=spatial_order)

    def c_struct(self):
        m = _lib.fb_map()
        m.P = self.P
        b = self._buf
        m.means3D, m.opacities = b["means3D"].data_ptr(), b["opacities"].data_ptr()
        m.scales, m.rotations, m.colors = b["scales"].data_ptr(), b["rotations"].data_ptr(), b["colors"].data_ptr()
        return m

    def tensors(self):
        return [self.means3D, self.colors, self.rotations, self.opacities, self.scales]


class FisherScorer:
    """Scores candidate poses by expected information gain, sum(H_candidate * 1/(H_train + lambda))."""

    deferred = True       # score_candidates / compute_H_train take sync=False (ShardedScorer.pose_eval: one sync per step)

    def __init__(self, cam: GaussianRasterizationSettings, C_H=4, grad_scale=1e-3, views_per_chunk=32,
                 inst_capacity=None, chunk_lanes=2, vis_capacity=None):
        """chunk_lanes=2: the workspace holds two chunks and `fb_score_candidates` alternates consecutive chunks
        between the caller's stream and a second stream (latency-bound binning kernels of one chunk run under the
        issue-bound kernels of the other); 1 = one chunk at a time, half the workspace."""
        assert C_H in (4, 11)
        assert chunk_lanes in (1, 2)
        self.chunk_lanes = chunk_lanes
        self.cam, self.C_H, self.grad_scale = cam, C_H, float(grad_scale)
        self.V = int(views_per_chunk)
        self.inst_capacity = inst_capacity
        self.vis_capacity = vis_capacity
        self._ws = None
        self.last_status = None
        self.last_instances = 0
        self.last_visible = 0
        self.launches = 0

    def _c_cam(self):
        rs = self.cam
        return _camera(rs.bg, rs.scale_modifier, rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, rs.image_height,
                       rs.image_width, rs.sh_degree, 0, rs.campos, rs.prefiltered)

    def _workspace(self, gmap, nbytes):
        """One grow-only device buffer shared by the scoring and the H_train entry points."""
        dev = gmap.means3D.device
        if self._ws is None or self._ws.device != dev or self._ws.numel() < nbytes:
            self._ws = None
            self._ws = torch.empty((int(nbytes),), dtype=torch.uint8, device=dev)
        return self._ws

    def _capacity(self, gmap):
        if self.inst_capacity is None:
            # about one (Gaussian, tile) instance per Gaussian and view in indoor maps; grow on overflow
            self.inst_capacity = max(1 << 20, int(1.5 * gmap.P) * self.V)
        return int(self.inst_capacity)

    def _vis_capacity(self, gmap):
        """Visible Gaussians (radii > 0) a single view may have; the per-view records are sized by it.  Small maps:
        everything; large maps: half the map (an indoor view sees 10-35 % of it), grown on overflow."""
        if self.vis_capacity is None or self.vis_capacity > gmap.P:
            self.vis_capacity = gmap.P if gmap.P <= (1 << 17) else max(1 << 17, gmap.P // 2)
        return int(self.vis_capacity)

    def _run(self, fn_name, gmap, w2c, extra, sync=True):
        """Common driver with grow-on-overflow.  sync=True reads the 32-byte status once per call (the only host
        sync) and re-runs with larger capacities if the device reported an overflow.  sync=False launches once and
        returns the device status tensor instead: the caller reads it together with its own results (one sync per
        planning step instead of one per call) and hands it to `accept_status`."""
        L = _lib.lib()
        dev = gmap.means3D.device
        w2c = _f32c(w2c.to(device=dev, dtype=torch.float32).reshape(-1, 4, 4))
        K = w2c.shape[0]
        cam, m = self._c_cam(), gmap.c_struct()
        status = torch.zeros((8,), dtype=torch.int32, device=dev)
        W, H = self.cam.image_width, self.cam.image_height
        while True:
            cap, vcap = self._capacity(gmap), self._vis_capacity(gmap)
            if fn_name == "fb_score_candidates":
                n = L.fb_score_workspace_bytes(gmap.P, W, H, self.V, cap, vcap) * self.chunk_lanes
            else:
                n = L.fb_hessian_workspace_bytes(gmap.P, W, H, self.V, cap)
            ws = self._workspace(gmap, n)
            n0 = L.fb_launch_count()
            with torch.cuda.device(dev):              # library state (second lane's stream, attributes) is per device
                rc = extra(L, cam, m, w2c, K, status, ws, n, cap, vcap)
            _lib.check(rc, fn_name)
            self.launches += int(L.fb_launch_count() - n0)
            if not sync:
                return status
            if self.accept_status(status.tolist(), fn_name, gmap.P):
                return status

    def accept_status(self, st, fn_name="fb_score_candidates", P=None):
        """Digest a status block read from the device (8 ints).  True: the call's results are valid.  False: a capacity
        was too small; it has been grown and the call must be repeated.  Other errors raise."""
        st = [int(x) for x in st]
        self.last_status = st
        self.last_instances = st[3]   # sum of num_rendered over the poses of this call
        self.last_visible = st[4]     # sum of (radii > 0) over the poses of this call
        if st[0] == 0:
            return True
        if st[0] == _lib.FB_E_CAPACITY:
            if st[1] > int(self.inst_capacity or 0):
                self.inst_capacity = int(st[1] * 1.25) + 1024
            if fn_name == "fb_score_candidates" and st[5] > int(self.vis_capacity or 0):
                self.vis_capacity = int(st[5] * 1.25) + 1024 if P is None else min(P, int(st[5] * 1.25) + 1024)
            return False
        _lib.check(st[0], fn_name)

    @torch.no_grad()
    def score_candidates(self, gmap: GaussianMap, w2c, H_train_inv, sync=True, map_order=False, weight_set=None):
        """scores[k] = sum(H_k * H_train_inv), H_k the Fisher diagonal of pose k (gaussian.py:1362-1370).
        w2c: [K,4,4] world-to-camera (already inverted).  Returns a CUDA float tensor [K].
        map_order=True: H_train_inv is already in the map's resident order (what compute_H_train(map_order=True)
        returns), no permutation.  sync=False: no host sync; returns (scores, status tensor), see `_run`.
        weight_set (i32 [K]): H_train_inv is [S,P,C_H] (map order) and pose k scores against tensor weight_set[k]."""
        K = w2c.reshape(-1, 4, 4).shape[0]
        dev = gmap.means3D.device
        if weight_set is not None:
            assert map_order, "weight sets are given in the map's resident order"
            weight = _f32c(H_train_inv.reshape(-1, gmap.P, self.C_H))
            wset = weight_set.to(device=dev, dtype=torch.int32).contiguous()
            assert wset.numel() == K
        else:
            H_train_inv = H_train_inv.reshape(gmap.P, self.C_H)
            weight = _f32c(H_train_inv if map_order else gmap.to_map_order(H_train_inv))
            wset = None
        scores = torch.zeros((K,), dtype=torch.float32, device=dev)

        def call(L, cam, m, w2c_t, K_, status, ws, nbytes, cap, vcap):
            return L.fb_score_candidates_sets(C.byref(cam), C.byref(m), w2c_t.data_ptr(), K_, weight.data_ptr(),
                                              None if wset is None else wset.data_ptr(), self.C_H,
                                              self.grad_scale, scores.data_ptr(), status.data_ptr(), ws.data_ptr(),
                                              nbytes, self.V, cap, vcap, _stream())
        status = self._run("fb_score_candidates", gmap, w2c, call, sync=sync)
        return scores if sync else (scores, status)

    @torch.no_grad()
    def compute_H_train(self, gmap: GaussianMap, keyframe_w2c, out=None, sync=True, map_order=False):
        """H_train[P,C_H] = sum over keyframes of the per-view Fisher diagonal (gaussian.py:1338-1348), in the
        caller's Gaussian order (map_order=True: in the map's resident order, no permutation).
        sync=False: no host sync; returns (H, status tensor), see `_run`."""
        H = torch.zeros((gmap.P, self.C_H), dtype=torch.float32, device=gmap.means3D.device)   # map order

        def call(L, cam, m, w2c_t, K_, status, ws, nbytes, cap, vcap):
            H.zero_()
            return L.fb_accumulate_hessian(C.byref(cam), C.byref(m), w2c_t.data_ptr(), K_, self.C_H, self.grad_scale,
                                           H.data_ptr(), status.data_ptr(), ws.data_ptr(), nbytes, self.V, cap,
                                           _stream())
        status = self._run("fb_accumulate_hessian", gmap, keyframe_w2c, call, sync=sync)
        if not map_order:
            H = gmap.to_caller_order(H)
        if out is not None:
            out += H
            H = out
        return H if sync else (H, status)

    @torch.no_grad()
    def compute_Hessian(self, gmap: GaussianMap, rel_w2c):
        """cur_H[P,C_H] for one pose (return_points=True form of the reference's compute_Hessian)."""
        return self.compute_H_train(gmap, rel_w2c.reshape(1, 4, 4))

    @torch.no_grad()
    def accumulate_hessian_sets(self, gmap: GaussianMap, w2c, H_sets, set_index):
        """H_sets[set_index[k]] += cur_H(pose k) for all k in one launch sequence (H_sets: [S,P,C_H], map order, updated in
        place): the `H_train_path = H_train_path + cur_H` of every path of a planning step at once."""
        dev = gmap.means3D.device
        assert H_sets.is_contiguous() and H_sets.shape[1:] == (gmap.P, self.C_H)
        sets = set_index.to(device=dev, dtype=torch.int32).contiguous()

        def call(L, cam, m, w2c_t, K_, status, ws, nbytes, cap, vcap):
            # (a repeat after a capacity overflow is safe: an overflowing chunk adds nothing)
            return L.fb_accumulate_hessian_sets(C.byref(cam), C.byref(m), w2c_t.data_ptr(), K_, self.C_H, self.grad_scale,
                                                H_sets.data_ptr(), sets.data_ptr(), status.data_ptr(), ws.data_ptr(), nbytes,
                                                self.V, cap, _stream())
        self._run("fb_accumulate_hessian", gmap, w2c, call)
        return H_sets

    @torch.no_grad()
    def plan_best_path(self, gmap: GaussianMap, paths_w2c, final_EIGs, H_train, H_reg_lambda=0.1, acc_H_train_every=5,
                       path_pose_weight=0.2, path_point_weight=1.0, path_end_weight=1.0, max_paths=21,
                       object_variant=False):
        """The path loop of plan_best_path (tester_gaussians_navigation.py:1664-1721) / path_evaluation (:1899-1973) for
        all candidate paths at once.  paths_w2c: list of [L_p,4,4] world-to-camera poses, one per action of path p (the
        caller's compute_next_campos + np.linalg.inv, :1686-1687); final_EIGs: the global goals' scores; H_train [P,C_H]
        in the caller's order.  Reference arithmetic, per path (only the first 21 paths are looked at, :1669-1672):
            step i = 1..L:  pose_EIG = log det(pose_H) with pose_H = I (compute_Hessian returns eye(6), gaussian.py:1570)
                            total += path_pose_weight * pose_EIG
                            if (i + 1) % acc_H_train_every == 0:  total += path_point_weight * log sum(cur_H / (H_path + lambda));
                                                                  H_path += cur_H
            total = total / L + path_end_weight * final_EIG   if path_end_weight > 0 else (total + final_EIG) / L
        (object_variant: point_EIG is overwritten with 0 and the end rule is total + object_path_end_weight * final_EIG,
        :1929-1960; pass that weight as path_end_weight.)  The point term of a non-accumulating step is computed and
        dropped by the reference; here it is never computed.  All paths' j-th accumulation poses are ONE
        fb_score_candidates_sets + ONE fb_accumulate_hessian_sets launch sequence (every path scores against, and adds
        to, its own running Hessian): L/acc levels instead of sum(L_p) per-pose Hessians with a host sync each.
        Returns (total_path_EIGs [n] list of floats, best index or None, point_EIG [n, levels] CPU tensor, NaN-padded)."""
        dev = gmap.means3D.device
        paths = [torch.as_tensor(p, dtype=torch.float32).reshape(-1, 4, 4) for p in list(paths_w2c)[:max_paths]]
        n = len(paths)
        acc = int(acc_H_train_every)
        steps = [[i for i in range(1, int(p.shape[0]) + 1) if (i + 1) % acc == 0] for p in paths]     # 1-based, :1703
        levels = max((len(s) for s in steps), default=0)
        point = torch.full((n, max(levels, 1)), float("nan"), dtype=torch.float32)
        totals = [0.0] * n
        if not object_variant and levels > 0 and n > 0:
            S = gmap.to_map_order(H_train.reshape(gmap.P, self.C_H)).unsqueeze(0).repeat(n, 1, 1).contiguous()
            for j in range(levels):
                active = [p for p in range(n) if len(steps[p]) > j]
                poses = torch.stack([paths[p][steps[p][j] - 1] for p in active]).to(dev)
                sets = torch.tensor(active, dtype=torch.int32, device=dev)
                W = torch.reciprocal(S + H_reg_lambda)
                sc = self.score_candidates(gmap, poses, W, map_order=True, weight_set=sets)
                e = torch.log(sc).cpu()                               # -inf if nothing is visible, as the reference
                for a, p in enumerate(active):
                    point[p, j] = e[a]
                    totals[p] += path_point_weight * float(e[a])
                if j + 1 < levels:
                    self.accumulate_hessian_sets(gmap, poses, S, sets)
        out = []
        for p in range(n):
            L = int(paths[p].shape[0])
            total = totals[p] + path_pose_weight * 0.0 * L           # log det(eye(6)) = 0 at every step
            fe = float(final_EIGs[p])
            if object_variant:
                total = total + path_end_weight * fe if path_end_weight > 0 else (total + fe) / max(L, 1)
            else:
                total = total / L + path_end_weight * fe if path_end_weight > 0 else (total + fe) / L
            out.append(total)
        best, best_val = None, -1.0                                   # best_path_EIG = -1., :1652 / :1891
        for p, v in enumerate(out):
            if v > best_val:
                best, best_val = p, v
        return out, best, point

    @torch.no_grad()
    def path_point_eig(self, gmap: GaussianMap, path_w2c, H_train, H_reg_lambda=0.1, acc_H_train_every=5):
        """Point-EIG terms along ONE candidate path, with the running Hessian of path_evaluation
        (tester_gaussians_navigation.py:1886-1975): for step i = 1..L (pose path_w2c[i-1])
            point_EIG_i = log sum(cur_H_i * 1/(H_path + lambda))      (0 if nothing is visible, :1933-1935)
            if (i + 1) % acc_H_train_every == 0:  total += point_EIG_i;  H_path += cur_H_i            (:1944-1946)
        Returns (point_EIG [L] CUDA tensor, total, H_path).  The weights only change at the accumulation steps, so
        every run of steps between two of them is ONE batched fb_score_candidates call and only the accumulation
        poses go through fb_accumulate_hessian: L/acc Hessians instead of L, no [P,C_H] tensor for the other poses.
        (The customised reference overwrites point_EIG with 0 after computing it, :1929/:1936, and adds a pose term
        log det(pose_H) from compute_Hessian(return_pose=True); this is the upstream FisherRF point term)"""
        path_w2c = path_w2c.reshape(-1, 4, 4)
        L = path_w2c.shape[0]
        acc = int(acc_H_train_every)
        H_path = H_train.clone()
        out = torch.zeros((L,), dtype=torch.float32, device=gmap.means3D.device)
        total = torch.zeros((), dtype=torch.float32, device=gmap.means3D.device)
        start = 0
        while start < L:
            b = start
            while b < L and (b + 2) % acc != 0:        # 0-based step b is an accumulation step iff (b + 2) % acc == 0
                b += 1
            last = min(b, L - 1)
            s = self.score_candidates(gmap, path_w2c[start:last + 1], torch.reciprocal(H_path + H_reg_lambda))
            e = torch.where(s > 0, torch.log(s.clamp_min(1e-45)), torch.zeros_like(s))
            out[start:last + 1] = e
            if b < L:
                total = total + e[-1]
                H_path = H_path + self.compute_Hessian(gmap, path_w2c[b])
            start = last + 1
        return out, total, H_path

    @torch.no_grad()
    def pose_eval(self, gmap: GaussianMap, poses_c2w, keyframe_w2c, H_reg_lambda=0.1):
        """Reference pose_eval: returns (scores [K] CPU tensor, poses)."""
        H_train = self.compute_H_train(gmap, keyframe_w2c)
        H_train_inv = torch.reciprocal(H_train + H_reg_lambda)
        poses_c2w = torch.as_tensor(poses_c2w, dtype=torch.float32, device=gmap.means3D.device).reshape(-1, 4, 4)
        w2c = torch.linalg.inv(poses_c2w)
        return self.score_candidates(gmap, w2c, H_train_inv).cpu(), poses_c2w

    # ---- POp-GS variants (gaussian_object.py:1545-1658, 2046-2109): Hutchinson probes through the same operator ----
    @torch.no_grad()
    def estimate_diag_JtJ(self, gmap: GaussianMap, w2c, K=4, probes=None, generator=None):
        """estimate_diag_JtJ_simple (gaussian_object.py:2046-2109): one forward at the pose, K backward passes of the
        power-2 operator with probe images z_k ([3,H,W], standard normal unless `probes` gives them), and
            diag = 1/K * sum_k g_k * g_k,   g_k = cat(d means3D [3P], d opacities [P], d rotations [4P], d scales [3P]).
        The forward is shared by the K probes (the reference keeps the graph, :2092); no autograd graph is built here.
        Returns (diag [11P] in the CALLER's Gaussian order, number of Gaussians with radius > 0)."""
        from .rasterizer import _C
        rs, dev = self.cam, gmap.means3D.device
        w2c = torch.as_tensor(w2c, dtype=torch.float32, device=dev).reshape(4, 4)
        pts4 = torch.cat([gmap.means3D, torch.ones((gmap.P, 1), dtype=torch.float32, device=dev)], dim=1)
        pts = (w2c @ pts4.T).T[:, :3].contiguous()                       # camera frame, :2062-2064
        e = torch.Tensor([])
        opac = gmap.opacities.reshape(-1, 1)
        R, im, radii, geom, inst, img, _ = _C.rasterize_gaussians(
            rs.bg, pts, gmap.colors, opac, gmap.scales, gmap.rotations, rs.scale_modifier, e, rs.viewmatrix,
            rs.projmatrix, rs.tanfovx, rs.tanfovy, rs.image_height, rs.image_width, e, rs.sh_degree, rs.campos,
            rs.prefiltered)
        acc = torch.zeros((gmap.P, 11), dtype=torch.float32, device=dev)
        for k in range(int(K)):
            z = probes[k] if probes is not None else torch.randn(im.shape, generator=generator, device=dev)
            _, _, g_op, g_mean, _, _, g_sc, g_rot = _C.rasterize_gaussians_backward(
                rs.bg, pts, radii, gmap.colors, gmap.scales, gmap.rotations, rs.scale_modifier, e, rs.viewmatrix,
                rs.projmatrix, rs.tanfovx, rs.tanfovy, z, e, rs.sh_degree, rs.campos, geom, R, inst, img, 2)
            g = torch.cat([g_mean, g_op, g_rot, g_sc], dim=1)
            acc.addcmul_(g, g)
            self.launches += 1
        acc = gmap.to_caller_order(acc / float(K))
        diag = torch.cat([acc[:, 0:3].reshape(-1), acc[:, 3].reshape(-1), acc[:, 4:8].reshape(-1),
                          acc[:, 8:11].reshape(-1)])                      # the reference's concatenation order, :2099
        return diag, int((radii > 0).sum().item())

    @staticmethod
    def topt_score_from_diags(H_train_diag, JtJ_diag, lam=1e-6):
        """T-optimality, to maximise: -sum 1/(H_train + JtJ + lam)   (gaussian_object.py:1705-1710)."""
        return -torch.sum(1.0 / torch.clamp(H_train_diag + JtJ_diag + lam, min=1e-12))

    @staticmethod
    def dopt_score_from_diags(H_train_diag, JtJ_diag, lam=1e-6):
        """D-optimality, to maximise: sum log(H_train + JtJ + lam) - sum log(H_train + lam)   (:1712-1718)."""
        Hm = H_train_diag + lam
        return torch.sum(torch.log(torch.clamp(Hm + JtJ_diag, min=1e-12))) - torch.sum(torch.log(torch.clamp(Hm, min=1e-12)))

    @torch.no_grad()
    def compute_H_train_popgs(self, gmap: GaussianMap, keyframe_w2c, K=4, generator=None):
        """Sum of the Hutchinson diagonals over the keyframes (gaussian_object.py:1545-1570)."""
        keyframe_w2c = torch.as_tensor(keyframe_w2c, dtype=torch.float32).reshape(-1, 4, 4)
        if keyframe_w2c.shape[0] == 0:
            raise RuntimeError("No keyframes available for POP-GS prior.")            # :1569
        H = None
        for w2c in keyframe_w2c:
            d, _ = self.estimate_diag_JtJ(gmap, w2c, K=K, generator=generator)
            H = d if H is None else H + d
        return H

    @torch.no_grad()
    def pose_eval_popgs(self, gmap: GaussianMap, poses_c2w, keyframe_w2c, criterion="topt", K=4, lam=1e-6,
                        generator=None):
        """pose_eval_popgs (gaussian_object.py:1619-1658): returns (scores [n] CPU tensor, poses)."""
        crit = criterion.lower()
        if crit not in ("topt", "dopt"):
            raise ValueError("criterion must be 'topt' or 'dopt'")                    # :1648
        score_fn = self.topt_score_from_diags if crit == "topt" else self.dopt_score_from_diags
        dev = gmap.means3D.device
        H_train = self.compute_H_train_popgs(gmap, keyframe_w2c, K=K, generator=generator)
        poses_c2w = torch.as_tensor(poses_c2w, dtype=torch.float32, device=dev).reshape(-1, 4, 4)
        scores = []
        for c2w in poses_c2w:
            cur, _ = self.estimate_diag_JtJ(gmap, torch.linalg.inv(c2w), K=K, generator=generator)
            scores.append(score_fn(H_train, cur, lam=lam))
        return torch.stack(scores).cpu(), poses_c2w

    # ---- POp-GS block form (gaussian_object.py:1572-1586, 1660-1703, 1720-1732, 2112-2176) ----
    @torch.no_grad()
    @_on_tensor_device
    def estimate_block_JtJ(self, gmap: GaussianMap, w2c, K=2, use_rot=True, use_scale=True, use_opacity=True,
                           probes=None, generator=None):
        """estimate_block_JtJ (gaussian_object.py:2112-2176): d x d blocks 1/K * sum_k g_k g_k^T of the Gaussians with
        radius > 0, g_k = [d means3D | d opacity | d rotations | d scales] under probe z_k through the power-2 operator.
        One forward, K backward launches, the outer products accumulated on device by `fb_block_outer_accumulate`
        (the block array is the only per-probe read-modify-write; no [N,d] concatenation, no autograd graph).
        Returns (H_blocks [Nv,d,d], vis_idx [Nv] int64 ascending, in the CALLER's Gaussian order)."""
        from .rasterizer import _C
        L = _lib.lib()
        rs, dev = self.cam, gmap.means3D.device
        w2c = torch.as_tensor(w2c, dtype=torch.float32, device=dev).reshape(4, 4)
        pts4 = torch.cat([gmap.means3D, torch.ones((gmap.P, 1), dtype=torch.float32, device=dev)], dim=1)
        pts = (w2c @ pts4.T).T[:, :3].contiguous()                       # camera frame, :2125-2127
        e = torch.Tensor([])
        opac = gmap.opacities.reshape(-1, 1)
        R, im, radii, geom, inst, img, _ = _C.rasterize_gaussians(
            rs.bg, pts, gmap.colors, opac, gmap.scales, gmap.rotations, rs.scale_modifier, e, rs.viewmatrix,
            rs.projmatrix, rs.tanfovx, rs.tanfovy, rs.image_height, rs.image_width, e, rs.sh_degree, rs.campos,
            rs.prefiltered)
        vis_idx = torch.where(gmap.to_caller_order(radii) > 0)[0]         # :2147, caller's indices, ascending
        if gmap.perm is None:
            rows = vis_idx
        else:                                                             # resident row of each visible Gaussian
            rows = gmap.inverse_perm().index_select(0, vis_idx)
        rows = rows.contiguous()
        Nv = int(vis_idx.numel())
        d = 3 + (1 if use_opacity else 0) + (4 if use_rot else 0) + (3 if use_scale else 0)
        H_blocks = torch.zeros((Nv, d, d), dtype=torch.float32, device=dev)
        K = int(K)
        for k in range(K):
            z = probes[k] if probes is not None else torch.randn(im.shape, generator=generator, device=dev)
            _, _, g_op, g_mean, _, _, g_sc, g_rot = _C.rasterize_gaussians_backward(
                rs.bg, pts, radii, gmap.colors, gmap.scales, gmap.rotations, rs.scale_modifier, e, rs.viewmatrix,
                rs.projmatrix, rs.tanfovx, rs.tanfovy, z, e, rs.sh_degree, rs.campos, geom, R, inst, img, 2)
            rc = L.fb_block_outer_accumulate(
                Nv, rows.data_ptr(), g_mean.data_ptr(), g_op.data_ptr() if use_opacity else None,
                g_rot.data_ptr() if use_rot else None, g_sc.data_ptr() if use_scale else None,
                float(K) if k == K - 1 else 0.0, H_blocks.data_ptr(), _stream())
            _lib.check(rc, "fb_block_outer_accumulate")
            self.launches += 2
        return H_blocks, vis_idx

    @torch.no_grad()
    def compute_H_train_blocks(self, gmap: GaussianMap, keyframe_w2c, K=2, generator=None, **kw):
        """compute_H_train_blocks (gaussian_object.py:1572-1586), including its alignment rule: the running sum and
        its index list are truncated to the smaller visible count, position by position."""
        keyframe_w2c = torch.as_tensor(keyframe_w2c, dtype=torch.float32).reshape(-1, 4, 4)
        Hm, vis_ref = None, None
        for w2c in keyframe_w2c:
            Hb, vis_idx = self.estimate_block_JtJ(gmap, w2c, K=K, generator=generator, **kw)
            if Hm is None:
                Hm, vis_ref = Hb, vis_idx
            else:
                Nv = min(Hm.shape[0], Hb.shape[0])
                Hm = Hm[:Nv] + Hb[:Nv]
                vis_ref = vis_ref[:Nv]
        if Hm is None:
            raise RuntimeError("No keyframes available for POP-GS prior (blocks).")   # :1585
        return Hm, vis_ref

    @_on_tensor_device
    def _block_opt(self, Hm_blocks, J_blocks, lam, criterion, idx_m=None, idx_j=None, per_block=False):
        L = _lib.lib()
        Hm_blocks, J_blocks = _f32c(Hm_blocks), _f32c(J_blocks)
        d = int(Hm_blocks.shape[-1])
        if J_blocks.shape[-1] != d or Hm_blocks.shape[-2] != d:
            raise ValueError("block shapes differ")
        n = int(idx_m.numel()) if idx_m is not None else int(Hm_blocks.shape[0])
        if idx_m is None and J_blocks.shape[0] != n:
            raise ValueError("Hm_blocks and J_blocks differ in length")
        dev = Hm_blocks.device
        partial = torch.zeros((max(1, L.fb_block_opt_partials(n)),), dtype=torch.float64, device=dev)
        terms = torch.empty((n,), dtype=torch.float32, device=dev) if per_block else None
        ptr = lambda t: None if t is None else t.contiguous().data_ptr()
        rc = L.fb_block_opt(n, d, Hm_blocks.data_ptr(), ptr(idx_m), J_blocks.data_ptr(), ptr(idx_j), float(lam),
                            criterion, partial.data_ptr(), ptr(terms), _stream())
        _lib.check(rc, "fb_block_opt")
        self.launches += 1
        total = partial.sum().to(torch.float32)
        return (total, terms) if per_block else total

    def t_opt_blocks(self, Hm_blocks, J_blocks, lam=1e-6):
        """T-optimality over blocks, to maximise: -sum_b trace((Hm_b + J_b + lam I)^-1)   (gaussian_object.py:1720-1724)."""
        return self._block_opt(Hm_blocks, J_blocks, lam, 0)

    def d_opt_blocks(self, Hm_blocks, J_blocks, lam=1e-6):
        """D-optimality over blocks: sum_b logdet(Hm_b + lam I + J_b) - logdet(Hm_b + lam I)   (:1726-1732)."""
        return self._block_opt(Hm_blocks, J_blocks, lam, 1)

    @torch.no_grad()
    def pose_eval_popgs_blocks(self, gmap: GaussianMap, poses_c2w, keyframe_w2c, criterion="topt", K=6, lam=1e-6,
                               use_rot=True, use_scale=True, use_opacity=True, generator=None):
        """pose_eval_popgs_blocks (gaussian_object.py:1660-1703): the blocks of the Gaussians visible both in the
        prior and from the candidate are paired by index (`np.intersect1d(..., return_indices=True)`, done here with a
        device-side searchsorted: both lists are ascending and unique) and gathered inside `fb_block_opt`.
        A candidate that shares no Gaussian with the prior scores -inf (:1682-1685).  Returns (scores CPU, poses)."""
        crit = criterion.lower()
        if crit not in ("topt", "dopt"):
            raise ValueError("criterion must be 'topt' or 'dopt'")                    # :1694
        kw = dict(use_rot=use_rot, use_scale=use_scale, use_opacity=use_opacity)
        dev = gmap.means3D.device
        Hm_blocks, train_vis_idx = self.compute_H_train_blocks(gmap, keyframe_w2c, K=K, generator=generator, **kw)
        poses_c2w = torch.as_tensor(poses_c2w, dtype=torch.float32, device=dev).reshape(-1, 4, 4)
        scores = []
        for c2w in poses_c2w:
            Jb, cur_vis_idx = self.estimate_block_JtJ(gmap, torch.linalg.inv(c2w), K=K, generator=generator, **kw)
            scores.append(self.score_blocks(Hm_blocks, train_vis_idx, Jb, cur_vis_idx, crit, lam))
        return torch.stack(scores).cpu(), poses_c2w

    def score_blocks(self, Hm_blocks, train_vis_idx, J_blocks, cur_vis_idx, criterion="topt", lam=1e-6):
        """The body of pose_eval_popgs_blocks' loop (:1671-1698) for one candidate: pair the prior's and the candidate's
        blocks by Gaussian index, score the pairs; -inf when no Gaussian is shared."""
        idx_train, idx_cur = intersect_sorted(train_vis_idx, cur_vis_idx)
        if idx_train.numel() == 0:
            return torch.tensor(float("-inf"), device=Hm_blocks.device)
        return self._block_opt(Hm_blocks, J_blocks, lam, 0 if criterion.lower() == "topt" else 1, idx_train, idx_cur)


def intersect_sorted(a, b):
    """Positions (ia, ib) with a[ia] == b[ib] for two ascending, duplicate-free int64 tensors: what
    `np.intersect1d(a, b, return_indices=True)[1:]` returns (gaussian_object.py:1674-1676), on the device."""
    if a.numel() == 0 or b.numel() == 0:
        z = torch.empty((0,), dtype=torch.int64, device=a.device)
        return z, z
    pos = torch.searchsorted(b, a).clamp_(max=b.numel() - 1)
    hit = b[pos] == a
    return torch.nonzero(hit).reshape(-1), pos[hit]
