"""Randomised soak on the GPU box: many regions of very different shapes, distance-matrix scores
against the UNMODIFIED reference (oracle/_ref, all host threads), plus gap-affine and edit CIGARs
against the reference aligners.  Not a test (minutes of CPU); prints mismatch counts."""
import ctypes as C, os, sys, time
sys.path.insert(0, os.getcwd()); sys.path.insert(0, os.path.join(os.getcwd(), "tests"))
import numpy as np
import support as S
import otter_b200 as ob
from otter_b200.capi import PAIR_DTYPE

seed = int(sys.argv[1]) if len(sys.argv) > 1 else 1
rng = np.random.default_rng(seed)
R = S.ref()
ctx = ob.Context(0)
vp = C.c_void_p
regions = []
for _ in range(int(sys.argv[2]) if len(sys.argv) > 2 else 400):
    if os.environ.get("SOAK_LONG"):  # long-read tier: 2048-slot rings, pairs finished on HBM rings
        tl = int(rng.choice([9000, 15000, 25000, 45000]))
        e = float(rng.choice([0.005, 0.05, 0.08]))
        n = int(rng.integers(2, 8))
    else:
        tl = int(rng.choice([20, 80, 300, 1000, 2500, 5000, 9000]))
        e = float(rng.choice([0.0, 0.002, 0.01, 0.05, 0.12, 0.3]))
        n = int(rng.integers(2, 26))
    reads = S.tr_region(rng, n, tl, e, extra_frac=float(rng.choice([0.0, 0.2, 0.6])), n_rate=float(rng.choice([0, 0, 0.1, 0.5])))
    if rng.random() < 0.1:
        reads[int(rng.integers(0, n))] = reads[0].copy()       # identical pair
    if rng.random() < 0.05:
        reads[int(rng.integers(0, n))] = np.frombuffer(b"N", dtype=np.uint8)
    regions.append(reads)
B = S.region_batch(regions)
t0 = time.time()
got = ctx.region_edit_scores(B["bases"], B["off"], B["lens"], B["region_start"], B["score_off"], B["n_scores"])
t1 = time.time()
lens = B["lens"]
# reference: distances (double) per region -> compare as score / max(len)
R.ref_fill_dist_matrix_regions.restype = C.c_double
R.ref_fill_dist_matrix_regions.argtypes = [vp, vp, vp, vp, C.c_uint32, vp, vp, C.c_int]
want = np.empty(B["n_scores"], dtype=np.float64)
R.ref_fill_dist_matrix_regions(S.ptr(B["bases"]), S.ptr(B["off"]), S.ptr(lens), S.ptr(B["region_start"]),
                               B["n_regions"], S.ptr(want), S.ptr(B["score_off"]), os.cpu_count())
t2 = time.time()
mine = np.empty_like(want)
q = 0
rs = B["region_start"]
for g in range(B["n_regions"]):
    L = lens[rs[g]:rs[g + 1]].astype(np.float64)
    n = len(L)
    iu = np.triu_indices(n, 1)
    m = np.maximum(L[iu[0]], L[iu[1]])
    k = len(m)
    mine[q:q + k] = got[q:q + k] / m
    q += k
bad = np.nonzero(mine.view(np.uint64) != want.view(np.uint64))[0]
print(f"seed {seed}: {B['n_regions']} regions, {B['n_scores']} pairs: gpu {t1-t0:.2f}s ref {t2-t1:.2f}s mismatches {len(bad)}", ctx.stats())
# CIGAR paths on a subset: every read against the first read of its region
rows = []
for g in range(0, B["n_regions"], 3):
    a = int(rs[g])
    for r in range(a + 1, int(rs[g + 1])):
        p, t = (a, r) if lens[a] >= lens[r] else (r, a)
        d = int(lens[p]) - int(lens[t])
        rows.append((p, t, -1, 0, 0, 0) if rng.random() < 0.6 else (p, t, d // 2, d - d // 2, 0, 0))
pairs = np.array(rows, dtype=PAIR_DTYPE)
req = np.array(rows, dtype=np.int32)
for edit, fn in ((False, R.ref_affine_cigars_batch), (True, R.ref_edit_cigars_batch)):
    sc, ln, begin, packed = ctx.cigars_packed(B["bases"], B["off"], lens, pairs, edit=edit)
    fn.restype = C.c_double
    clen = np.zeros(len(rows), dtype=np.int32)
    fn(S.ptr(B["bases"]), S.ptr(B["off"]), S.ptr(lens), S.ptr(req), C.c_uint64(len(rows)), os.cpu_count(), S.ptr(clen))
    print(f"  {'edit' if edit else 'affine'} CIGAR lengths: {len(rows)} alignments, mismatches {(clen != ln).sum()}, retried {ctx.stats()['n_pairs_wide']}")
