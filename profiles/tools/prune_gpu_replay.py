"""prune_gpu_replay.py <dir with the out*.txt of oracle/tools/prune_search.c>: every pair on which the
exact prune fired in the CPU search (scores in the record headers, confirmed against the unmodified
reference by oracle/tools/replay_prune_search.py) through the GPU kernels, every ring tier."""
import glob, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import otter_b200
from otter_b200.capi import PAIR_DTYPE
recs = []
for f in sorted(glob.glob(os.path.join(sys.argv[1], "out*.txt"))):
    rows = open(f).read().split("\n")
    for i in range(0, len(rows) - 2, 3):
        h = rows[i].split()
        recs.append((int(h[4]), rows[i + 1].encode(), rows[i + 2].encode()))
print("pairs", len(recs), flush=True)
reads, want = [], []
for sc, p, t in recs:
    reads += [np.frombuffer(p, np.uint8), np.frombuffer(t, np.uint8)]
    want.append(sc)
want = np.array(want, dtype=np.int32)
lens = np.array([len(r) for r in reads], dtype=np.uint32)
off = np.zeros(len(reads), dtype=np.uint64); off[1:] = np.cumsum(lens[:-1], dtype=np.uint64)
bases = np.concatenate(reads)
pairs = np.zeros(len(recs), dtype=PAIR_DTYPE)
pairs["pattern"] = np.arange(len(recs)) * 2; pairs["text"] = np.arange(len(recs)) * 2 + 1; pairs["pbf"] = -1
for name, env in (("default tiers", {}), ("2048-slot ring forced", {"OTGPU_RING_SLOTS": "2048"}),
                  ("1024-slot ring forced + HBM pass", {"OTGPU_RING_SLOTS": "1024"}),
                  ("2048-slot ring, second kernel for the overflow", {"OTGPU_RING_SLOTS": "2048", "OTGPU_WIDE_TWO_PASS": "1"})):
    for k in ("OTGPU_RING_SLOTS", "OTGPU_WIDE_TWO_PASS"): os.environ.pop(k, None)
    os.environ.update(env)
    with otter_b200.Context(0) as c:
        t0 = time.perf_counter()
        got = c.edit_scores(bases, off, lens, pairs)
        dt = time.perf_counter() - t0
        print(f"{name}: {int((got != want).sum())} of {len(want)} scores differ from the reference's; pairs on the HBM ring {c.stats()['n_pairs_wide']}; {dt*1e3:.0f} ms", flush=True)
