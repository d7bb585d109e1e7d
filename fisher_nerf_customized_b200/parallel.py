"""Candidate-pose sharding over the GPUs of one box (one process per GPU, torch.distributed).

The reference has no multi-GPU path; this is the additive data-parallel layer of
SURVEY 8(e): every rank holds a replica of the Gaussian map; candidates are split into contiguous
blocks of ceil(K/G); keyframes for H_train are dealt round-robin and combined with one all-reduce;
per-candidate scores are all-gathered so that every rank computes the same argmax.  The only
collectives are the ones the path really has (map broadcast, H_train sum, score gather): there is
no data-path exchange inside a view.
"""
import torch
import torch.distributed as dist


def shard_bounds(K, world, rank):
    """Contiguous block [lo, hi) of candidate indices owned by `rank`."""
    per = (K + world - 1) // world if K > 0 else 0
    lo = min(K, rank * per)
    return lo, min(K, lo + per)


def round_robin(n, world, rank):
    return list(range(rank, n, world))


class ShardedScorer:
    """Wraps a local scorer (FisherScorer-like: .score_candidates, .compute_H_train) with the collectives.

    `group` may be any torch.distributed backend: nccl over NVLink on the GPU box, gloo in the CPU tests
    (where the local scorer is a stand-in; the product's local scorer is always the CUDA one)."""

    def __init__(self, scorer, group=None):
        self.scorer = scorer
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0

    def broadcast_map(self, tensors, src=0):
        """Map update: the owner's parameter tensors overwrite every replica (56*P bytes at init;
        callers pass only appended rows in the active-mapping loop)."""
        if self.world > 1:
            for t in tensors:
                dist.broadcast(t, src=src, group=self.group)
        return tensors

    def broadcast_appended(self, store, r0=None, n=None, src=0):
        """Delta map update: the rows rank `src` appended to its `ResidentParams` store (the
        `(r0, n)` that `append_from_frame` returned there) are written behind the live rows of every replica:
        a 16-byte header, then ONE broadcast of 56 (+4 timestep) bytes per NEW Gaussian instead of 56*P for the whole map.
        Every replica must hold the same live rows before the call.  Returns (r0, n) on every rank."""
        dev = store.timestep.device
        hdr = torch.tensor([int(r0 or 0), int(n or 0)], dtype=torch.int64, device=dev)
        if self.world > 1:
            dist.broadcast(hdr, src=src, group=self.group)
        r0, n = (int(x) for x in hdr.tolist())
        if self.rank != src:
            if store.n != r0:
                raise RuntimeError(f"replica out of step: {store.n} live rows here, owner appended at row {r0}")
            store._grow(r0 + n)
        if self.world > 1 and n > 0:
            # one payload collective: the new rows of the five arrays + timestep packed side by side ([n, 15] floats;
            # 13 if isotropic), instead of six launch-latency-bound broadcasts of row slices
            keys = sorted(store.buf)
            cols = [store.buf[k].shape[1] for k in keys]
            pack = torch.empty((n, sum(cols) + 1), dtype=torch.float32, device=dev)
            if self.rank == src:
                torch.cat([store.buf[k][r0:r0 + n] for k in keys] + [store.timestep[r0:r0 + n, None]], dim=1, out=pack)
            dist.broadcast(pack, src=src, group=self.group)
            if self.rank != src:
                o = 0
                for k, c in zip(keys, cols):
                    store.buf[k][r0:r0 + n] = pack[:, o:o + c]
                    o += c
                store.timestep[r0:r0 + n] = pack[:, o]
        if self.rank != src:
            store.n = r0 + n
            if hasattr(store, "_notify"):
                store._notify(r0, n)                      # attached GaussianMaps follow (the owner's did in append_from_frame)
        return r0, n

    def compute_H_train(self, gmap, keyframe_w2c):
        n = keyframe_w2c.shape[0]
        mine = round_robin(n, self.world, self.rank)
        H = None
        if len(mine) > 0:
            H = self.scorer.compute_H_train(gmap, keyframe_w2c[mine])
        if H is None:
            H = torch.zeros((gmap.P, self.scorer.C_H), dtype=torch.float32, device=gmap.means3D.device)
        if self.world > 1:
            dist.all_reduce(H, op=dist.ReduceOp.SUM, group=self.group)
        return H

    def score_candidates(self, gmap, w2c, H_train_inv):
        """Returns the full scores[K] on every rank."""
        K = w2c.shape[0]
        if K == 0:
            return torch.zeros((0,), dtype=torch.float32, device=gmap.means3D.device)
        lo, hi = shard_bounds(K, self.world, self.rank)
        per = (K + self.world - 1) // self.world if K > 0 else 0
        dev = gmap.means3D.device
        local = torch.zeros((per,), dtype=torch.float32, device=dev)
        if hi > lo:
            local[: hi - lo] = self.scorer.score_candidates(gmap, w2c[lo:hi], H_train_inv)
        if self.world == 1:
            return local[:K]
        gathered = torch.empty((self.world * per,), dtype=torch.float32, device=dev)
        dist.all_gather_into_tensor(gathered, local, group=self.group) if dev.type == "cuda" else \
            self._gather_cpu(gathered, local)
        return gathered[:K]

    def _gather_cpu(self, out, local):
        parts = [torch.empty_like(local) for _ in range(self.world)]
        dist.all_gather(parts, local, group=self.group)
        out.copy_(torch.cat(parts))

    def pose_eval(self, gmap, poses_c2w, keyframe_w2c, H_reg_lambda=0.1):
        """One planning step (pose_eval, gaussian.py:1354-1375) over the ranks: H_train from round-robin keyframes +
        all-reduce, reciprocal, this rank's block of candidates, all-gather, argmax.  Returns (scores [K], argmax) --
        identical on every rank.
        With a scorer that can defer its status check (FisherScorer: `deferred`), the whole step has ONE host sync:
        both status blocks travel with the scores in the all-gather payload and are read with them; if any rank reports
        a capacity overflow every rank grows and the step is repeated (all ranks see all status blocks, so they agree)."""
        if not getattr(self.scorer, "deferred", False):
            H_train = self.compute_H_train(gmap, keyframe_w2c)
            H_train_inv = torch.reciprocal(H_train + H_reg_lambda)
            w2c = torch.linalg.inv(poses_c2w.float())
            scores = self.score_candidates(gmap, w2c, H_train_inv)
            return scores, (int(torch.argmax(scores).item()) if scores.numel() else -1)
        dev = gmap.means3D.device
        sc = self.scorer
        K = int(poses_c2w.shape[0])
        w2c = torch.linalg.inv(poses_c2w.to(device=dev, dtype=torch.float32).reshape(-1, 4, 4))
        kf = keyframe_w2c.to(device=dev, dtype=torch.float32).reshape(-1, 4, 4)
        mine = round_robin(int(kf.shape[0]), self.world, self.rank)
        lo, hi = shard_bounds(K, self.world, self.rank)
        per = (K + self.world - 1) // self.world if K > 0 else 0
        zero_status = torch.zeros((8,), dtype=torch.int32, device=dev)
        while True:
            if mine:
                H, st_h = sc.compute_H_train(gmap, kf[mine], sync=False, map_order=True)
            else:
                H, st_h = torch.zeros((gmap.P, sc.C_H), dtype=torch.float32, device=dev), zero_status
            if self.world > 1:
                dist.all_reduce(H, op=dist.ReduceOp.SUM, group=self.group)
            H_inv = torch.reciprocal(H + H_reg_lambda)
            payload = torch.zeros((per + 16,), dtype=torch.int32, device=dev)
            st_s = zero_status
            if hi > lo:
                local, st_s = sc.score_candidates(gmap, w2c[lo:hi], H_inv, sync=False, map_order=True)
                payload[: hi - lo] = local.view(torch.int32)
            payload[per: per + 8] = st_h
            payload[per + 8: per + 16] = st_s
            if self.world > 1:
                gathered = torch.empty((self.world * (per + 16),), dtype=torch.int32, device=dev)
                if dev.type == "cuda":
                    dist.all_gather_into_tensor(gathered, payload, group=self.group)
                else:
                    self._gather_cpu(gathered, payload)
            else:
                gathered = payload
            host = gathered.cpu().reshape(self.world, per + 16)          # the step's one host sync
            ok = True
            for r in range(self.world):
                # every rank digests every rank's status: capacities grow alike, so replicas stay in step
                mine_r = r == self.rank
                st_h_r, st_s_r = host[r, per: per + 8].tolist(), host[r, per + 8: per + 16].tolist()
                if mine_r:
                    ok_h = sc.accept_status(st_h_r, "fb_accumulate_hessian", gmap.P)
                    ok_s = sc.accept_status(st_s_r, "fb_score_candidates", gmap.P)
                    ok = ok and ok_h and ok_s
                else:
                    ok = ok and st_h_r[0] == 0 and st_s_r[0] == 0
            if ok:
                scores = host[:, :per].contiguous().view(torch.float32).reshape(-1)[:K]
                return scores, (int(torch.argmax(scores)) if K else -1)
