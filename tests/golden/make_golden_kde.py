#!/usr/bin/env python
"""Generate tests/golden/kde_golden.json from the UNMODIFIED reference (oracle/_ref): KDE::f
(src/ankde.cpp:19-23) on the 0.0025 grid of otter_find_clustering_dist for a few distance sets,
values and densities stored as hex floats (the comparison is bit for bit), and whole
otter_hclust results (labels, ic, fc) for distance matrices of 1..60 reads.
Run in the build container:   python tests/golden/make_golden_kde.py"""
import ctypes as C
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import support as S  # noqa: E402
from test_gpu_kde import distances, grid_points  # noqa: E402


def ref_hclust(R, dist, lens, max_alleles):
    n = len(lens)
    labels = np.zeros(n, dtype=np.int32)
    ic, fc = C.c_int(), C.c_int()
    R.ref_otter_hclust.restype = None
    R.ref_otter_hclust.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_int, C.c_void_p,
                                   C.POINTER(C.c_int), C.POINTER(C.c_int)]
    R.ref_otter_hclust(S.ptr(dist), n, S.ptr(lens), max_alleles, S.ptr(labels), C.byref(ic), C.byref(fc))
    return labels.tolist(), ic.value, fc.value


def cluster_case(rng, n, kind):
    """A condensed distance matrix of n reads from 1-3 alleles with within/between distances."""
    k = 1 + kind % 3
    lab = rng.integers(0, k, n)
    within = [0.01, 0.04, 0.08][kind % 3]
    between = [0.1, 0.2, 0.35][(kind // 3) % 3]
    d = []
    for i in range(n):
        for j in range(i + 1, n):
            base = within if lab[i] == lab[j] else between * (1 + abs(int(lab[i]) - int(lab[j])) * 0.5)
            d.append(abs(rng.normal(base, base * 0.25)))
    lens = rng.integers(200, 900, n).astype(np.uint32) if kind % 2 else rng.integers(600, 4000, n).astype(np.uint32)
    return np.minimum(np.array(d, dtype=np.float64), 1.0), lens


def main():
    R = S.ref()
    R.ref_kde_densities.restype = C.c_int
    R.ref_kde_densities.argtypes = [C.c_void_p, C.c_uint64, C.c_double, C.c_double, C.c_void_p, C.c_int]
    rng = np.random.default_rng(0xDE)
    grid = grid_points()
    cases = []
    for g in range(8):
        v = distances(rng, int(rng.integers(3, 20)), g % 4)
        h = 0.01 if g % 2 else 0.015
        out = np.zeros(len(grid) + 8)
        k = R.ref_kde_densities(S.ptr(v), len(v), h, 0.0025, S.ptr(out), len(out))
        assert k == len(grid)
        cases.append(dict(h=h, values=[float(x).hex() for x in v], densities=[float(x).hex() for x in out[:k]]))
    clusters = []
    for t in range(36):
        n = [1, 2, 3, 4][t] if t < 4 else int(rng.integers(3, 33))
        dist, lens = cluster_case(rng, n, t)
        for ma in ((2,) if t % 5 else (2, 1, 0, 3)):
            labels, ic, fc = ref_hclust(R, dist, lens, ma)
            clusters.append(dict(n=n, max_alleles=ma, dist=[float(x).hex() for x in dist], lens=lens.tolist(),
                                 labels=labels, ic=ic, fc=fc))
    path = os.path.join(HERE, "kde_golden.json")
    with open(path, "w") as f:
        json.dump(dict(source="oracle/_ref (unmodified reference KDE::f / otter_hclust)", cases=cases,
                       clusters=clusters), f)
    import collections
    print(path, os.path.getsize(path), "bytes", collections.Counter(c["fc"] for c in clusters))


if __name__ == "__main__":
    main()
