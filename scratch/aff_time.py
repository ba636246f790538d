import time, sys, os, numpy as np
sys.path.insert(0, os.getcwd())
import bench, otter_b200
from otter_b200.capi import PAIR_DTYPE, _p
class A: workload="C2"; regions=200
B = bench.workload_arrays("C2", 0, 200)
req = bench.consensus_requests(B)
n = len(req)
pairs = np.zeros(n, dtype=PAIR_DTYPE)
for i, f in enumerate(("pattern", "text", "pbf", "pef", "tbf", "tef")):
    pairs[f] = req[:, i]
ctx = otter_b200.Context(0)
lens, off, bases = B["lens"], B["off"], B["bases"]
for it in range(3):
    t0 = time.perf_counter()
    caps = lens[pairs["pattern"]].astype(np.uint64) + lens[pairs["text"]].astype(np.uint64)
    coff = np.zeros(n + 1, dtype=np.uint64); coff[1:] = np.cumsum(caps)
    scores = np.zeros(n, dtype=np.int32); clen = np.zeros(n, dtype=np.int32)
    cig = np.zeros(int(coff[-1]) + 1, dtype=np.uint8)
    t1 = time.perf_counter()
    ctx._ck(ctx._l.otgpu_affine_cigars(ctx._h, _p(bases), bases.size, _p(off), _p(lens), len(lens), _p(pairs), n, 4, 6, 2, _p(scores), _p(cig), _p(coff), _p(clen)))
    t2 = time.perf_counter()
    out = [cig[int(coff[i]):int(coff[i]) + int(clen[i])].tobytes() for i in range(n)]
    t3 = time.perf_counter()
    ctx._ck(ctx._l.otgpu_affine_cigars(ctx._h, _p(bases), bases.size, _p(off), _p(lens), len(lens), _p(pairs), n, 4, 6, 2, _p(scores), _p(cig), _p(coff), _p(clen)))
    t4 = time.perf_counter()
    ctx._ck(ctx._l.otgpu_affine_cigars(ctx._h, _p(bases), bases.size, _p(off), _p(lens), len(lens), _p(pairs), n, 4, 6, 2, _p(scores), None, None, _p(clen)))
    t5 = time.perf_counter()
    print("prep %.1f  call(fresh buf) %.1f  slice %.1f  call(warm buf) %.1f  call(lengths only) %.1f ms; cap bytes %d, cigar bytes %d" % (1e3*(t1-t0), 1e3*(t2-t1), 1e3*(t3-t2), 1e3*(t4-t3), 1e3*(t5-t4), coff[-1], clen.sum()))
