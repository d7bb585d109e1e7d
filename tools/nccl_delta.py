"""torchrun --nproc-per-node 2 tools/nccl_delta.py : delta map update over NCCL (owner densifies a frame, replicas
receive only the new rows), then both ranks score the same candidates on the grown map and compare."""
import os, sys
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, os.getcwd())
import fisher_nerf_customized_b200 as fb
from fisher_nerf_customized_b200 import synth
from tests.golden.make_golden import frame_inputs
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl")
H = W = 256
store = fb.ResidentParams(capacity=50_000)
sh = fb.ShardedScorer(None)
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
tot = 0.0
for step in range(8):
    r0 = n = None
    if rank == 0:
        color, depth, K, w2c, mask = frame_inputs(H, W, 100 + step)
        tc, td, tK, tw, tm = [torch.from_numpy(np.ascontiguousarray(x)).cuda() for x in (color, depth, K, w2c, mask)]
        r0, n = store.append_from_frame(tc, td, tK, tw, mask=tm, time_idx=step)
    torch.cuda.synchronize(); dist.barrier()
    a.record()
    r0, n = sh.broadcast_appended(store, r0, n, src=0)
    b.record(); b.synchronize()
    if step >= 2:
        tot += a.elapsed_time(b)
chk = torch.stack([v.double().sum() for _, v in sorted(store.params().items())] + [store.timestep[:store.n].double().sum()])
allc = [torch.zeros_like(chk) for _ in range(world)]
dist.all_gather(allc, chk)
same = all(torch.equal(allc[0], c) for c in allc)
if rank == 0:
    print(f"delta map update over NCCL, {world} ranks: {store.n} Gaussians after 8 frames (~{n} rows = {60 * n / 1e6:.2f} MB per update), "
          f"{tot / 6:.3f} ms per update, replicas identical: {same}")
dist.destroy_process_group()
