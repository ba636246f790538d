"""Seeded synthetic tandem-repeat read sets of the shapes BASELINE.json names.

Region: 100/500 bp random flanks around motif x copies; two alleles (allele 1 = allele 0 plus up
to 20 % extra copies); each read picks an allele and gets per-base sub/ins/del errors at rate e
(HiFi-like 0.005, ONT-like 0.05 where half of the inserted bases repeat the previous base);
1 % of reads carry one 'N'.  Pure numpy; the same bytes feed the CUDA path and the CPU oracle.
"""
import numpy as np

ALPHA = np.frombuffer(b"ACGT", dtype=np.uint8)
_CODE = np.zeros(256, dtype=np.int64)
_CODE[ALPHA] = np.arange(4)


def rand_seq(rng, n):
    return ALPHA[rng.integers(0, 4, n)]


def mutate(rng, seq, e, n_rate=0.0, homopolymer=False):
    n = len(seq)
    if n == 0 or e <= 0:
        out = seq.copy()
    else:
        hit = rng.random(n) < e
        nh = int(hit.sum())
        kind = np.zeros(n, dtype=np.int8)  # 0 keep, 1 sub, 2 ins (after base), 3 del
        kind[hit] = rng.integers(1, 4, nh)
        sub = kind == 1
        s2 = seq.copy()
        s2[sub] = ALPHA[(_CODE[seq[sub]] + rng.integers(1, 4, int(sub.sum()))) % 4]
        reps = np.ones(n, dtype=np.int64)
        reps[kind == 3] = 0
        reps[kind == 2] = 2
        out = np.repeat(s2, reps)
        ins_pos = (np.cumsum(reps) - 1)[kind == 2]
        new = ALPHA[rng.integers(0, 4, len(ins_pos))]
        if homopolymer and len(ins_pos):
            keep = rng.random(len(ins_pos)) < 0.5
            new = np.where(keep, out[ins_pos - 1], new)
        out[ins_pos] = new
    if n_rate > 0 and len(out) and rng.random() < n_rate:
        out[rng.integers(0, len(out))] = ord("N")
    return out


def tr_region(rng, n_reads, target_len, e, flank=100, extra_frac=0.2, n_rate=0.01, vntr_p=0.1,
              mixed=False):
    """One region's reads.  mixed=True: each read is HiFi-like or ONT-like with p=0.5 (C3)."""
    mlen = int(rng.integers(20, 61)) if rng.random() < vntr_p else int(rng.integers(2, 7))
    motif = rand_seq(rng, mlen)
    copies0 = max(1, (target_len - 2 * flank) // mlen)
    copies1 = copies0 + int(rng.integers(0, int(copies0 * extra_frac) + 1))
    fl, fr = rand_seq(rng, flank), rand_seq(rng, flank)
    alleles = [np.concatenate([fl, np.tile(motif, c), fr]) for c in (copies0, copies1)]
    reads = []
    for _ in range(n_reads):
        er = e if not mixed else (0.005 if rng.random() < 0.5 else 0.05)
        reads.append(mutate(rng, alleles[int(rng.integers(0, 2))], er, n_rate,
                            homopolymer=er >= 0.02))
    return reads


def pack_reads(reads):
    """list of uint8 arrays -> (bases u8, off u64, len u32)."""
    lens = np.array([len(r) for r in reads], dtype=np.uint32)
    off = np.zeros(len(reads), dtype=np.uint64)
    if len(reads):
        off[1:] = np.cumsum(lens[:-1], dtype=np.uint64)
    bases = (np.concatenate(reads).astype(np.uint8) if len(reads) and lens.sum()
             else np.zeros(0, np.uint8))
    return bases, off, lens


def region_batch(regions):
    """list of regions (each a list of reads) -> flat arrays for the region ABI."""
    reads = [r for reg in regions for r in reg]
    bases, off, lens = pack_reads(reads)
    rs = np.zeros(len(regions) + 1, dtype=np.uint32)
    rs[1:] = np.cumsum([len(r) for r in regions])
    n = np.diff(rs).astype(np.uint64)
    npairs = n * (n - (n > 0)) // 2
    so = np.zeros(len(regions), dtype=np.uint64)
    if len(regions):
        so[1:] = np.cumsum(npairs[:-1])
    return dict(bases=bases, off=off, lens=lens, region_start=rs, score_off=so,
                n_scores=int(npairs.sum()), reads=reads, n_regions=len(regions))


# BASELINE.json configs (index = position in "configs")
CONFIGS = {
    "C1": dict(n_regions=1000, n_reads=30, len_lo=2000, len_hi=5000, e=0.005, flank=100,
               mixed=False, desc="1,000 TR regions x 30 HiFi reads (2-5 kb, 0.5% error)"),
    "C2": dict(n_regions=1000, n_reads=60, len_lo=2000, len_hi=5000, e=0.05, flank=100,
               mixed=False, desc="1,000 TR regions x 60 ONT-like reads (2-5 kb, 5% error)"),
    "C3": dict(n_regions=200, n_reads=40, len_lo=10000, len_hi=20000, e=0.05, flank=500,
               mixed=True, desc="200 VNTR regions x 40 reads of 10-20 kb (half HiFi, half ONT-like)"),
    "C4": dict(n_regions=100000, n_reads=30, len_lo=2000, len_hi=5000, e=0.005, flank=100,
               mixed=False, desc="100k TR regions x 30 HiFi reads"),
}


def make_regions(name, seed=0, n_regions=None):
    c = CONFIGS[name]
    rng = np.random.default_rng(0xB200 + 1000 * int(name[1:]) + seed)
    nr = n_regions if n_regions is not None else c["n_regions"]
    out = []
    for _ in range(nr):
        tl = int(rng.integers(c["len_lo"], c["len_hi"] + 1))
        out.append(tr_region(rng, c["n_reads"], tl, c["e"], flank=c["flank"], mixed=c["mixed"]))
    return out


def pair_lengths(B):
    """(|x|,|y|) of every pair of a region batch in score order (for bytes / GCUPS accounting)."""
    lens = B["lens"].astype(np.int64)
    rs = B["region_start"]
    xs, ys = [], []
    for g in range(len(rs) - 1):
        l = lens[rs[g]:rs[g + 1]]
        i, j = np.triu_indices(len(l), 1)
        xs.append(l[i])
        ys.append(l[j])
    if not xs:
        return np.zeros(0, np.int64), np.zeros(0, np.int64)
    return np.concatenate(xs), np.concatenate(ys)


# ---- the 100 k-region HiFi description of BASELINE.json configs[3] / north_star ("C4S") -----------
# Every region has its own seed, so any rank can generate any region: one description, generated
# per shard on the fly, partitioned by estimated work, nothing shipped between ranks.
C4S_SEED = 0xB2004
C4S_REGIONS = 100_000


def c4s_target_len(g):
    return int(np.random.default_rng([C4S_SEED, int(g)]).integers(2000, 5001))


def c4s_region(g):
    rng = np.random.default_rng([C4S_SEED, int(g)])
    tl = int(rng.integers(2000, 5001))  # the first draw: c4s_target_len reproduces it
    return tr_region(rng, 30, tl, 0.005, flank=100)


def c4s_work(n_regions=C4S_REGIONS):
    """Estimated alignment work per region from its header alone (30 reads of about tl bases:
    pairs x |x| x |y| grows with tl^2), without generating a read."""
    tl = np.array([c4s_target_len(g) for g in range(n_regions)], dtype=np.float64)
    return tl * tl


def _c4s_pack(gs):
    lens, parts = [], []
    for g in gs:
        reg = c4s_region(g)
        lens += [len(r) for r in reg]
        parts += reg
    return (np.concatenate(parts).astype(np.uint8) if parts else np.zeros(0, np.uint8),
            np.array(lens, dtype=np.uint32), len(gs))


def c4s_batch(region_ids, procs=1):
    """Flat arrays (region ABI) of the listed regions of the description, in the listed order;
    generated by `procs` worker processes."""
    region_ids = [int(g) for g in region_ids]
    if procs > 1 and len(region_ids) >= 4 * procs:
        import multiprocessing as mp
        per = max(50, (len(region_ids) + 8 * procs - 1) // (8 * procs))
        chunks = [region_ids[i:i + per] for i in range(0, len(region_ids), per)]
        try:  # (workers only run numpy; a stuck pool must not take the caller down with it)
            with mp.get_context("fork").Pool(procs) as pool:
                outs = pool.map_async(_c4s_pack, chunks).get(timeout=600)
        except Exception:
            outs = [_c4s_pack(c) for c in chunks]
    else:
        outs = [_c4s_pack(region_ids)]
    bases = np.concatenate([o[0] for o in outs]) if outs else np.zeros(0, np.uint8)
    lens = np.concatenate([o[1] for o in outs]) if outs else np.zeros(0, np.uint32)
    n_reg = len(region_ids)
    off = np.zeros(len(lens), dtype=np.uint64)
    if len(lens):
        off[1:] = np.cumsum(lens[:-1], dtype=np.uint64)
    rs = np.arange(n_reg + 1, dtype=np.uint32) * 30
    npairs = np.full(n_reg, 435, dtype=np.uint64)
    so = np.zeros(n_reg, dtype=np.uint64)
    if n_reg:
        so[1:] = np.cumsum(npairs[:-1])
    return dict(bases=bases, off=off, lens=lens, region_start=rs, score_off=so,
                n_scores=int(npairs.sum()), n_regions=n_reg)
