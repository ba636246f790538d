# cost of the 8-bit (non-ACGT) pairs: same batch with and without its N bases, end-to-end wall time
import sys, time, numpy as np
sys.path.insert(0, ".")
import otter_b200
from otter_b200 import synth
for name, nreg in (("C1", 1000), ("C2", 300)):
    B = synth.region_batch(synth.make_regions(name, 0, nreg))
    clean = B["bases"].copy(); clean[clean == ord("N")] = ord("A")
    with otter_b200.Context(0) as ctx:
        for tag, bases in (("with N", B["bases"]), ("N->A", clean)):
            ts = []
            for _ in range(6):
                t = time.perf_counter()
                ctx.region_edit_scores(bases, B["off"], B["lens"], B["region_start"], B["score_off"], B["n_scores"])
                ts.append(time.perf_counter() - t)
            st = ctx.stats()
            print(name, tag, "ms %.2f" % (1e3 * min(ts)), "pairs", B["n_scores"], "8bit", st["n_pairs_8bit"], flush=True)
