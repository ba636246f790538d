"""N>1 path on CPU: two gloo ranks shard one batch region-wise, each computes its shard (with the
oracle port standing in for the GPU, which is not present here), and a host-side gather rebuilds
the full score vector.  No data-path collective: only the final gather uses the process group."""
import os
import sys

import numpy as np
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import support as S  # noqa: E402
from harness import sharding  # noqa: E402


def make_batch():
    rng = np.random.default_rng(4)
    regions = [S.tr_region(rng, int(rng.integers(1, 7)), int(rng.integers(100, 500)), 0.05,
                           extra_frac=0.4) for _ in range(11)]
    return S.region_batch(regions)


def port_region_scores(B):
    P = S.port()
    out = np.zeros(B["n_scores"], dtype=np.int32)
    rs = B["region_start"]
    q = 0
    for g in range(B["n_regions"]):
        for i in range(rs[g], rs[g + 1]):
            for j in range(i + 1, rs[g + 1]):
                a, b = (i, j) if B["lens"][i] >= B["lens"][j] else (j, i)
                x = S.b(B["bases"][int(B["off"][a]):int(B["off"][a]) + int(B["lens"][a])])
                y = S.b(B["bases"][int(B["off"][b]):int(B["off"][b]) + int(B["lens"][b])])
                out[q] = P.wfo_edit_score(x, len(x), y, len(y), 0, 0, 0, 0, 0, None)
                q += 1
    return out


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    B = make_batch()
    work = sharding.estimate_region_work(B["lens"], B["region_start"])
    shards = sharding.partition_regions(work, world)
    mine = sharding.shard_batch(B, shards[rank])
    scores = port_region_scores(mine)
    gathered = [None] * world
    dist.all_gather_object(gathered, (shards[rank].tolist(), scores.tolist()))
    if rank == 0:
        full = np.full(B["n_scores"], -99, dtype=np.int32)
        for idx, sc in gathered:
            sharding.scatter_scores(B, idx, np.array(sc, dtype=np.int32), full)
        q.put(full.tolist())
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_region_sharding_gloo():
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    full = np.array(q.get(timeout=120), dtype=np.int32)
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    B = make_batch()
    assert np.array_equal(full, port_region_scores(B))


def test_partition_covers_and_balances():
    rng = np.random.default_rng(0)
    work = rng.random(1000) ** 4
    for world in (1, 2, 4, 8):
        shards = sharding.partition_regions(work, world)
        allr = np.sort(np.concatenate(shards))
        assert np.array_equal(allr, np.arange(1000))
        loads = np.array([work[s].sum() for s in shards])
        assert loads.max() <= 1.05 * loads.mean() + work.max()
