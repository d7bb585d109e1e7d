"""CPU, world_size 2 over gloo: the candidate sharding / H_train reduction / score gather logic of
ShardedScorer with a stand-in local scorer (the product's local scorer is always CUDA)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from fisher_nerf_customized_b200.parallel import ShardedScorer, round_robin, shard_bounds


class FakeMap:
    def __init__(self, P):
        self.P = P
        self.means3D = torch.zeros((P, 3))


class FakeScorer:
    """score(pose) = sum(weight) * pose[0,3]; Hessian(pose) = pose[0,3] * ones: linear, so sums are checkable."""
    C_H = 4

    def score_candidates(self, gmap, w2c, weight):
        return weight.sum() * w2c[:, 0, 3]

    def compute_H_train(self, gmap, kf):
        return kf[:, 0, 3].sum() * torch.ones((gmap.P, self.C_H))


class FakeDeferredScorer(FakeScorer):
    """The sync=False protocol of FisherScorer: results + a device status block, digested by accept_status.  The first
    scoring call of rank 1 reports a capacity overflow, so every rank must repeat the step."""
    deferred = True

    def __init__(self, rank):
        self.rank, self.calls, self.grown = rank, 0, 0

    def compute_H_train(self, gmap, kf, sync=True, map_order=False):
        assert sync is False and map_order is True
        return FakeScorer.compute_H_train(self, gmap, kf), torch.zeros(8, dtype=torch.int32)

    def score_candidates(self, gmap, w2c, weight, sync=True, map_order=False):
        assert sync is False and map_order is True
        st = torch.zeros(8, dtype=torch.int32)
        self.calls += 1
        if self.rank == 1 and self.calls == 1:
            st[0], st[1] = -4, 12345
        return FakeScorer.score_candidates(self, gmap, w2c, weight), st

    def accept_status(self, st, fn_name="", P=None):
        if st[0] == 0:
            return True
        assert st[0] == -4
        self.grown += 1
        return False


def _poses(n, off=0.0):
    p = torch.eye(4).repeat(n, 1, 1)
    p[:, 0, 3] = torch.arange(n, dtype=torch.float32) + 1 + off
    return p


def _worker(rank, world, port, K, NKF, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        gmap = FakeMap(10)
        sh = ShardedScorer(FakeScorer())
        H = sh.compute_H_train(gmap, _poses(NKF))
        w = torch.full((10, 4), 0.5)
        s = sh.score_candidates(gmap, _poses(K, 0.25), w)
        t = [torch.full((3,), float(rank))]
        sh.broadcast_map(t, src=0)
        scores, best = sh.pose_eval(gmap, torch.linalg.inv(_poses(K, 0.25)), _poses(NKF), 0.1)
        # the one-sync form of the same step (status blocks ride in the all-gather payload; rank 1 overflows once)
        dsc = FakeDeferredScorer(rank)
        scores_d, best_d = ShardedScorer(dsc).pose_eval(gmap, torch.linalg.inv(_poses(K, 0.25)), _poses(NKF), 0.1)
        assert torch.equal(scores_d, scores) and best_d == best
        if K >= 2:
            assert dsc.calls == 2 and dsc.grown == (1 if rank == 1 else 0)    # every rank repeated the step
        out[rank] = (H.numpy(), s.numpy(), t[0].numpy(), scores.numpy(), best)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("K,NKF", [(7, 3), (2, 1), (1, 5), (0, 2)])
def test_sharded_scorer_world2(K, NKF):
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(2, port, K, NKF, out), nprocs=2, join=True)
    H_want = float(sum(range(1, NKF + 1)))
    s_want = 20.0 * (np.arange(K, dtype=np.float32) + 1.25)
    for r in (0, 1):
        H, sc, t, scores, best = out[r]
        np.testing.assert_allclose(H, H_want * np.ones((10, 4)))
        np.testing.assert_allclose(sc, s_want, rtol=1e-6)
        assert sc.shape == (K,)
        np.testing.assert_array_equal(t, np.zeros(3))          # rank 0's map overwrote the replica
        if K:
            assert best == K - 1
    if K:
        np.testing.assert_array_equal(out[0][3], out[1][3])    # identical scores (and argmax) on every rank


def test_partition_helpers():
    for K in (0, 1, 5, 8, 256, 257):
        for world in (1, 2, 3, 8):
            got = []
            for r in range(world):
                lo, hi = shard_bounds(K, world, r)
                assert 0 <= lo <= hi <= K
                got += list(range(lo, hi))
            assert got == list(range(K))
            rr = sorted(sum((round_robin(K, world, r) for r in range(world)), []))
            assert rr == list(range(K))


def _append_worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from fisher_nerf_customized_b200.mapupdate import ResidentParams
        store = ResidentParams(capacity=4, device="cpu")          # small: the second update forces every replica to grow
        sh = ShardedScorer(FakeScorer())
        g = torch.Generator().manual_seed(7)
        for step, n_new in enumerate((3, 5, 0)):
            r0 = n = None
            if rank == 0:                                          # the owner "densifies": rows written behind the live ones
                r0, n = store.n, n_new
                store._grow(r0 + n)
                for k, v in store.buf.items():
                    v[r0:r0 + n] = torch.rand((n, v.shape[1]), generator=g) + step
                store.timestep[r0:r0 + n] = float(step)
                store.n = r0 + n
            got = sh.broadcast_appended(store, r0, n, src=0)
            assert got == (sum((3, 5, 0)[:step]), n_new)
        out[rank] = ({k: v.clone().numpy() for k, v in store.params().items()}, store.timestep[:store.n].numpy().copy(), store.n)
        if rank == 1:                                              # a replica that missed an update refuses the next one
            store.n -= 1
        try:
            sh.broadcast_appended(store, store.n if rank == 0 else None, 0, src=0)
            out[rank] += (None,)
        except RuntimeError as e:
            out[rank] += (str(e),)
    finally:
        dist.destroy_process_group()


def test_delta_map_update_world2():
    """ShardedScorer.broadcast_appended: only the appended rows travel; replicas grow in step and end up identical."""
    world = 2
    with socket.socket() as sk:
        sk.bind(("127.0.0.1", 0))
        port = sk.getsockname()[1]
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_append_worker, args=(world, port, out), nprocs=world, join=True)
    p0, t0, n0, e0 = out[0]
    p1, t1, n1, e1 = out[1]
    assert n0 == n1 == 8
    for k in p0:
        assert p0[k].shape[0] == 8 and np.array_equal(p0[k], p1[k]), k
    assert np.array_equal(t0, t1) and t0.tolist() == [0, 0, 0, 1, 1, 1, 1, 1]
    assert e0 is None and "out of step" in e1
