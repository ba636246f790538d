"""Region sharding for multi-GPU runs: regions are independent, so the batch is
partitioned region-wise by ESTIMATED WORK (not by count) and each rank aligns its shard on its
own GPU; the only cross-rank step is a host-side gather of the int32 scores.  No collective on
the data path."""
import numpy as np


def estimate_region_work(lens, region_start, err_hint=0.05):
    """~ sum over pairs of (length difference + 2*err*length)^2 — WFA cost grows with score^2."""
    lens = np.asarray(lens, dtype=np.float64)
    w = np.zeros(len(region_start) - 1)
    for g in range(len(w)):
        l = lens[region_start[g]:region_start[g + 1]]
        if len(l) < 2:
            continue
        i, j = np.triu_indices(len(l), 1)
        s = np.abs(l[i] - l[j]) + 2.0 * err_hint * np.maximum(l[i], l[j])
        w[g] = float((s * s).sum())
    return w


def partition_regions(work, world):
    """Greedy longest-processing-time partition; returns `world` sorted index arrays."""
    order = np.argsort(-np.asarray(work), kind="stable")
    load = np.zeros(world)
    shards = [[] for _ in range(world)]
    for g in order:
        r = int(np.argmin(load))
        shards[r].append(int(g))
        load[r] += work[g]
    return [np.array(sorted(s), dtype=np.int64) for s in shards]


def shard_batch(B, regions_idx):
    """Sub-batch (flat arrays for the region ABI) holding only the listed regions."""
    rs = B["region_start"]
    keep_reads = np.concatenate([np.arange(rs[g], rs[g + 1]) for g in regions_idx]) \
        if len(regions_idx) else np.zeros(0, np.int64)
    lens = B["lens"][keep_reads]
    off = np.zeros(len(lens), dtype=np.uint64)
    if len(lens):
        off[1:] = np.cumsum(lens[:-1], dtype=np.uint64)
    bases = np.concatenate([B["bases"][int(B["off"][r]):int(B["off"][r]) + int(B["lens"][r])]
                            for r in keep_reads]) if len(keep_reads) else np.zeros(0, np.uint8)
    n = np.array([rs[g + 1] - rs[g] for g in regions_idx], dtype=np.uint64)
    region_start = np.zeros(len(regions_idx) + 1, dtype=np.uint32)
    region_start[1:] = np.cumsum(n)
    npairs = n * (n - (n > 0)) // 2
    score_off = np.zeros(len(regions_idx), dtype=np.uint64)
    if len(regions_idx):
        score_off[1:] = np.cumsum(npairs[:-1])
    return dict(bases=bases, off=off, lens=lens.astype(np.uint32), region_start=region_start,
                score_off=score_off, n_scores=int(npairs.sum()), n_regions=len(regions_idx))


def scatter_scores(B, regions_idx, shard_scores, out):
    """Place a shard's scores into the full batch's score array (host-side gather)."""
    rs, so = B["region_start"], B["score_off"]
    q = 0
    for g in regions_idx:
        n = int(rs[g + 1] - rs[g])
        k = n * (n - 1) // 2 if n else 0
        out[int(so[g]):int(so[g]) + k] = shard_scores[q:q + k]
        q += k
    return out
