import sys, time
import os; sys.path.insert(0, os.getcwd()); sys.path.insert(0, os.path.join(os.getcwd(),'tests'))
import numpy as np
from support import *
import otter_b200 as ob
rng = np.random.default_rng(7)
P = port()
regions = []
for tl, e, nr in [(300,0.005,8),(300,0.05,8),(3500,0.005,10),(3500,0.05,10),(1000,0.3,6),(12000,0.05,6)]:
    for _ in range(3):
        regions.append(tr_region(rng, nr, tl, e, extra_frac=0.3))
regions.append([np.frombuffer(b"N",np.uint8), np.frombuffer(b"N",np.uint8), np.frombuffer(b"ACGTN",np.uint8), np.frombuffer(b"A",np.uint8)])
B = region_batch(regions)
ctx = ob.Context(0)
t0=time.time()
got = ctx.region_edit_scores(B['bases'], B['off'], B['lens'], B['region_start'], B['score_off'], B['n_scores'])
print("gpu time", time.time()-t0, ctx.stats())
exp = np.zeros_like(got)
q = 0
for reg in regions:
    for i in range(len(reg)):
        for j in range(i+1, len(reg)):
            x, y = b(reg[i]), b(reg[j])
            if len(x) < len(y): x, y = y, x
            exp[q] = P.wfo_edit_score(x, len(x), y, len(y), 0,0,0,0,0,None); q += 1
bad = np.nonzero(got != exp)[0]
print("pairs", len(got), "mismatches", len(bad))
for i in bad[:10]: print(i, got[i], exp[i])
# ends-free explicit pairs
reads = [r for reg in regions[:12] for r in reg]
bases, off, lens = pack_reads(reads)
pairs = []
for i in range(0, len(reads)-1):
    j = i+1
    p, t = (i, j) if lens[i] >= lens[j] else (j, i)
    d = int(lens[p]) - int(lens[t])
    for pb, pe in ((0,d),(d,0),(d//2,d//2)):
        pairs.append((p,t,pb,pe,0,0))
    pairs.append((p,t,-1,0,0,0))
pairs = np.array(pairs, dtype=ob.capi.PAIR_DTYPE)
got = ctx.edit_scores(bases, off, lens, pairs)
exp = np.array([P.wfo_edit_score(b(reads[p]), int(lens[p]), b(reads[t]), int(lens[t]), int(pb>=0), max(pb,0), pe, tb, te, None) for p,t,pb,pe,tb,te in pairs.tolist()], dtype=np.int32)
bad = np.nonzero(got != exp)[0]
print("ef pairs", len(got), "mismatches", len(bad), ctx.stats())
for i in bad[:10]: print(i, pairs[i], got[i], exp[i])
