#!/usr/bin/env python
"""Generate tests/golden/linkage_golden.json from hclust-cpp as vendored in the reference
(include/hclust-cpp, via oracle/_ref): hclust_fast(AVERAGE) + cutree_k + cutree_cdist on random
condensed matrices incl. heavy ties.  Run in the build container:
    python tests/golden/make_golden_linkage.py"""
import ctypes as C
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import support as S  # noqa: E402
from test_linkage import ARGS, matrix, run  # noqa: E402


def main():
    R = S.ref()
    R.ref_hclust_average.restype = None
    R.ref_hclust_average.argtypes = ARGS
    rng = np.random.default_rng(0x11C)
    cases = []
    for t in range(48):
        n = [2, 3, 4, 5][t] if t < 4 else int(rng.integers(3, 28))
        dist = matrix(rng, n, t % 4)
        k = int(rng.integers(0, n + 2))
        cut = float(rng.choice(dist)) if t % 2 else float(rng.random())
        merge, height, lk, lh = run(R.ref_hclust_average, dist, n, k, cut)
        cases.append(dict(n=n, k=k, cut=float(cut).hex(), dist=[float(x).hex() for x in dist], merge=merge,
                          height=height, labels_k=lk, labels_h=lh))
    path = os.path.join(HERE, "linkage_golden.json")
    with open(path, "w") as f:
        json.dump(dict(source="oracle/_ref: hclust-cpp (fastcluster) as vendored in the reference", cases=cases), f)
    print(path, len(cases), "cases", os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
