"""CPU, world_size 2, gloo: the N>1 plumbing -- every rank derives the same shard cuts from the
map, owns a contiguous SNP range, and the host-side gather reassembles the per-SNP arrays in .bim
order with no data-path collective.  The compute itself needs a GPU, so each
rank fills its slice from the committed golden vectors."""
import os
import sys

import numpy as np
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from nldsc_b200 import capi
    z = np.load(os.path.join(ROOT, "tests", "golden", "n257_m1500_wide.npz"))
    pos, wind = z["positions"], float(z["ld_wind"])
    cuts = capi.plan_shards(pos, wind, world)
    lo, hi = capi.plan_windows(pos, wind)
    a, b = int(cuts[rank]), int(cuts[rank + 1])
    # rows this rank would read: its range widened by the halo, 64-aligned on the left
    load = (int(lo[a]) // 64 * 64, int(hi[b - 1]) + 1) if b > a else (0, 0)
    part = {k: z["ref64_" + k][a:b] for k in ("l2", "l2d", "l2_ws")}
    gathered = [None] * world
    dist.all_gather_object(gathered, (a, b, load, part))
    if rank == 0:
        full = {k: np.empty_like(z["ref64_" + k]) for k in part}
        for (ra, rb, _, p) in gathered:
            for k in p:
                full[k][ra:rb] = p[k]
        ok = all(np.array_equal(full[k], z["ref64_" + k], equal_nan=True) for k in full)
        ranges = [(g[0], g[1]) for g in gathered]
        halos = [g[2] for g in gathered]
        q.put((ok, ranges, halos))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_shard_and_gather():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    ok, ranges, halos = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert ok
    assert ranges[0][0] == 0 and ranges[0][1] == ranges[1][0] and ranges[1][1] == 1500
    assert halos[0][0] <= ranges[0][0] and halos[0][1] >= ranges[0][1]
    assert halos[1][0] <= ranges[1][0] and halos[1][1] == 1500
